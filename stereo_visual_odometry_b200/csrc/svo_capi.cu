// svo_capi.cu -- C-ABI host layer (include/svo.h) over the sm_100a kernels.
//
// Owns the device-resident state of VisualOdometry (include/vo.h:233-268 of the
// reference): frame id, the t0/t1 image pyramids, currentVOFeatures, the rotation /
// translation guess and last_transform, for n_seq independent sequences.  Every branch
// of stereo_callback (vo.cpp:41-137) is a device-side predicate, so the host never looks
// at intermediate results: a frame is a fixed set of launches on four streams --
//   stream_in  ingest (H2D, pyramid, FAST) of frame f + 1
//   stream_k   speculative circular matching of the candidates of frame f + 1
//   stream     the dependent chain of frame f: bucketing -> (left-over KLT) -> select ->
//              RANSAC hypotheses -> inliers / feature set
//   stream_fb  LM pose refit + result of frame f
// ordered by events.  For a single sequence the streams live in three green contexts
// (disjoint SM partitions), so the latency-bound pose chain never queues behind the
// throughput-bound tracking kernels.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only NVTX 3: ranges per stage (no-ops unless a tool is attached)

#include "svo_kernels.cuh"

using namespace svo;

#define SVO_RESULT_RING 32
#define SVO_EV_RING 32   // per-frame events live in rings of this many frames
#define SVO_EV_MASK (SVO_EV_RING - 1)
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                            \
            return (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver) ? SVO_ERR_NO_DEVICE \
                                                                                   : SVO_ERR_CUDA;  \
        }                                                                                          \
    } while (0)

extern "C" const char* svo_version(void) { return "svo-b200 0.1 (sm_100a)"; }
extern "C" const char* svo_last_error(void) { return g_err.c_str(); }
extern "C" int svo_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(SVO_ERR_NO_DEVICE, cudaGetErrorString(e));
    }
    return n;
}

extern "C" void svo_default_params(svo_params* p, int width, int height) {
    memset(p, 0, sizeof(*p));
    p->width = width;
    p->height = height;
    p->n_seq = 1;
    p->buckets_along_height = 92;
    p->buckets_along_width = 160;
    p->bucket_start_row = 4;
    p->features_per_bucket = 1;
    p->features_threshold = 15;
    p->pre_matching_threshold = 100;
    p->age_threshold = 20;
    p->fast_threshold = 20;
    p->win_size = 10;
    p->max_level = 3;
    p->lk_max_iters = 30;
    p->lk_epsilon = 1e-4f;
    p->lk_min_eig = 1e-3f;
    p->circular_threshold = 0.15f;
    p->ransac_reproj_error = 8.f;
    p->ransac_iterations = 100;
    p->ransac_hypotheses = 100;
    p->ransac_confidence = 0.98f;
    p->max_translation_norm = 0.1;
    p->max_rotation_norm = 0.5;
    p->ransac_seed = 0;   /* unused: the sample stream is OpenCV's fixed RNG((uint64)-1) */
    p->device = 0;
    p->channels = 1;
    p->lookahead = 0;
}

// ---------------------------------------------------------------------------
struct svo_ctx {
    svo_params p;
    PyrGeom pg;
    BucketGeom bg;
    PipeConfig pc;
    DevArrays d;
    cudaStream_t stream = nullptr;      // dependent chain (bucket .. inliers) and everything synchronous
    cudaStream_t stream_in = nullptr;   // ingest of the next frame (H2D, pyramid, FAST)
    cudaStream_t stream_k = nullptr;    // speculative KLT of the FAST-winner candidates (and the un-cached fallback)
    cudaStream_t stream_k2 = nullptr;   // speculative KLT of the matched-feature candidates (same partition)
    cudaEvent_t ev_spf[SVO_EV_RING] = {};   // FAST-winner candidates of frame f tracked
    cudaEvent_t ev_ts[SVO_EV_RING] = {};    // the tracked-feature pass of frame f is ready to start (its inputs are complete)
    cudaEvent_t ev_bm[SVO_EV_RING] = {};    // bucket merge of frame f done (predictions for f + 1)
    cudaStream_t stream_fb = nullptr;   // pose refit + result
    cudaStream_t stream_h2d = nullptr;  // host frames -> staging ring (copy engine only)
    cudaEvent_t ev_h2d[SVO_EV_RING] = {};   // staging copy of frame f complete
    cudaStream_t stream_full = nullptr; // whole-GPU stream for serial frames (== stream when there are no partitions)
    int klt_blocks_full = 0;            // persistent KLT blocks per sequence on the whole GPU
    void* green[3] = {nullptr, nullptr, nullptr};   // CUgreenCtx of the chain / ingest / KLT partitions
    int sm_chain = 0, sm_in = 0, sm_k = 0;          // SMs of each partition (0: no partitioning)
    bool spec = false;                  // speculative tracking on
    struct TraceRec { const char* name; long long frame; int stream; cudaEvent_t a, b; };
    std::vector<TraceRec> trace;        // SVO_TRACE=<file>: per-kernel timeline of frames 30..45, dumped by svo_destroy
    const char* trace_path = nullptr;
    bool pipeline_ctx = true;           // false: private helper context of a stateless call (plain streams, no speculation)
    int klt_blocks_k = 0;               // persistent KLT blocks per sequence on the KLT partition
    int f_share = 100;                  // % of them the FAST-candidate kernel is launched with (SVO_FSHARE).  Both speculative passes get the
                                        // full grid: T's stream has the higher priority and F(f + 1) may not start before T(f) is ready, so
                                        // T's blocks are placed first and F fills what T leaves (59 % 2749 / 2845, 70 % 2915 / 2980,
                                        // 85 % 3038 / 3056, 100 % 3094 / 3075 frames/s resident / e2e)
    int t_share = 100;                  // % of them the tracked-feature kernel is launched with (SVO_TSHARE; 0 = 100 - f_share): blocks
                                        // beyond its resident share start as the FAST-candidate kernel's blocks leave
    cudaEvent_t ev_in[SVO_EV_RING] = {};    // ingest(f) complete
    cudaEvent_t ev_td[SVO_EV_RING] = {};    // every bucketed feature of f has its circular-matching result
    cudaEvent_t ev_sel[SVO_EV_RING] = {};   // select(f) done: pyramids / keys of f - 1 are free
    cudaEvent_t ev_trk[SVO_EV_RING] = {};   // speculative KLT of frame f complete
    cudaEvent_t ev_fa[SVO_EV_RING] = {};    // chain of frame f complete
    cudaEvent_t ev_fb[SVO_EV_RING] = {};    // refit + result of frame f complete
    cudaEvent_t ev_done[SVO_RESULT_RING];                           // result of frame f in host_result ring
    std::vector<void*> allocs;
    uint8_t* stage_L = nullptr;
    uint8_t* stage_R = nullptr;
    svo_frame_result* host_result = nullptr;   // pinned, [SVO_RESULT_RING][n_seq]
    long long frames = 0;       // frames submitted
    long long collected = 0;    // frames whose result was handed to the caller (host API only)
    long long tracked = 0;      // frames whose result goes to the host ring (submitted through the host API)
    long long launches = 0;
    int last_collected = 0;     // ring index of the result returned by the last svo_collect
    bool have_projection = false;
    int klt_blocks = 148;   // persistent KLT blocks per sequence
    // timers
    bool timers = false;
    cudaEvent_t ev[7];
    double stage_ms[6] = {0, 0, 0, 0, 0, 0};
    long long stage_launches[6] = {0, 0, 0, 0, 0, 0};
};

static BucketGeom make_bucket_geom(int height, int width, int n_bh, int n_bw, int start_row, int age_thr,
                                   int fast_thr) {
    BucketGeom bg;
    bg.n_bh = n_bh;
    bg.n_bw = n_bw;
    bg.bh = (height + n_bh - 1) / n_bh;
    bg.bw = (width + n_bw - 1) / n_bw;
    bg.start_row = start_row;
    bg.age_threshold = age_thr;
    bg.fast_threshold = fast_thr;
    bg.nb = n_bh * n_bw;
    return bg;
}

// cv::buildOpticalFlowPyramid stops when a level would be <= winSize in either dimension
static int channels_of(const svo_params& p) { return p.channels == 3 ? 3 : 1; }
static PyrGeom make_pyr_geom(int w, int h, int max_level, int win, int cn = 1) {
    PyrGeom pg;
    memset(&pg, 0, sizeof(pg));
    pg.cn = cn;
    long long off = 0;
    int lw = w, lh = h, l = 0;
    for (; l <= max_level && l < SVO_MAX_LEVELS; ++l) {
        if (l > 0) {
            const int nw = (lw + 1) / 2, nh = (lh + 1) / 2;
            if (nw <= win || nh <= win) break;
            lw = nw;
            lh = nh;
        }
        pg.w[l] = lw;
        pg.h[l] = lh;
        pg.pitch[l] = (lw * cn + 127) & ~127;
        pg.off[l] = off;
        off += (long long)pg.pitch[l] * lh;
    }
    pg.nlevels = l;
    for (int k = l; k < SVO_MAX_LEVELS; ++k) {   // keep unused levels pointing at valid memory
        pg.w[k] = pg.w[l - 1];
        pg.h[k] = pg.h[l - 1];
        pg.pitch[k] = pg.pitch[l - 1];
        pg.off[k] = pg.off[l - 1];
    }
    pg.image_bytes = (off + 255) & ~255ll;
    return pg;
}

static LkConfig make_lk_config(const svo_params& p, int nlevels) {
    LkConfig c;
    c.nlevels = nlevels;
    c.max_iters = p.lk_max_iters < 0 ? 0 : (p.lk_max_iters > 100 ? 100 : p.lk_max_iters);
    double e = p.lk_epsilon < 0.f ? 0.0 : (p.lk_epsilon > 10.f ? 10.0 : (double)p.lk_epsilon);
    c.eps_sq = e * e;
    c.min_eig = p.lk_min_eig;
    lk_config_finish(c);
    return c;
}

template <class T>
static cudaError_t dev_alloc(svo_ctx* c, T** ptr, size_t count, bool zero = true) {
    void* q = nullptr;
    const size_t bytes = (count ? count : 1) * sizeof(T);
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e != cudaSuccess) return e;
    c->allocs.push_back(q);
    if (zero) e = cudaMemsetAsync(q, 0, bytes, c->stream);
    *ptr = (T*)q;
    return e;
}

static int validate_params(const svo_params* p) {
    if (!p) return fail(SVO_ERR_INVALID, "params is NULL");
    if (p->width < 16 || p->height < 16 || p->width > 8192 || p->height > 8192)
        return fail(SVO_ERR_INVALID, "image size out of range (16..8192)");
    if ((long long)p->width * p->height >= (1ll << 24)) return fail(SVO_ERR_INVALID, "image has >= 2^24 pixels");
    if (p->n_seq < 1) return fail(SVO_ERR_INVALID, "n_seq < 1");
    if (p->channels != 0 && p->channels != 1 && p->channels != 3)
        return fail(SVO_ERR_INVALID, "channels must be 1 (8UC1) or 3 (interleaved 8UC3, the reference CLI's images)");
    if (p->win_size != SVO_LK_WIN) return fail(SVO_ERR_INVALID, "only winSize 10x10 (vo.h:251) is built");
    if (p->max_level < 0 || p->max_level >= SVO_MAX_LEVELS) return fail(SVO_ERR_INVALID, "max_level must be 0..3");
    if (p->buckets_along_height < 1 || p->buckets_along_width < 1) return fail(SVO_ERR_INVALID, "bad bucket grid");
    if (p->fast_threshold < 4 || p->fast_threshold > 255) return fail(SVO_ERR_INVALID, "fast_threshold must be 4..255");
    if (p->lookahead < 0 || p->lookahead > 16) return fail(SVO_ERR_INVALID, "lookahead must be 0..16");
    if (p->lookahead > 1 && (p->n_seq != 1 || p->channels == 3))
        return fail(SVO_ERR_INVALID, "lookahead needs a single-sequence 8UC1 context");
    if (p->ransac_hypotheses < 1 || p->ransac_hypotheses > 4096)
        return fail(SVO_ERR_INVALID, "ransac_hypotheses must be 1..4096");
    return SVO_OK;
}

// Streams of a context.  A single-sequence context is latency bound: its pose chain (a few small blocks) must
// never wait for a block slot behind the persistent tracking kernels, so the SMs are split with CUDA green
// contexts into chain / ingest / KLT partitions (SVO_PARTITION="chain,ingest" SM counts, "0" = off, ingest 0 =
// the ingest shares the chain partition; default 16,0 of the 148 SMs: the rest tracks).  Batched contexts are throughput bound and use plain streams on the whole GPU.
// SVO_SPEC=0/1 overrides whether the speculative tracking runs (default: single-sequence contexts only).
static int create_streams(svo_ctx* c, const svo_params* p) {
    int want_chain = 8, want_in = 0;   // (measured, resident / e2e frames/s: 8 SMs 3186 / 3178, 16 SMs 3093 / 3107, 16 + 8 for the ingest 2986 / 2986)
    c->spec = p->n_seq == 1 && c->pipeline_ctx && channels_of(*p) == 1;   // 8UC3 mode: plain serial frames
    if (const char* e = getenv("SVO_SPEC"))
        if (c->pipeline_ctx && channels_of(*p) == 1) c->spec = atoi(e) != 0;
    if (const char* e = getenv("SVO_PARTITION")) {
        want_chain = want_in = 0;
        sscanf(e, "%d,%d", &want_chain, &want_in);
    } else if (getenv("NV_COMPUTE_PROFILER_PERFWORKS_DIR")) {
        // Nsight Compute cannot always set up kernels launched into green contexts ("Failed to prepare kernel for
        // profiling"): under ncu the context runs unpartitioned unless SVO_PARTITION asks for it explicitly
        want_chain = want_in = 0;
        static bool told = false;
        if (!told) fprintf(stderr, "svo: running under Nsight Compute, SM partitions off (set SVO_PARTITION to force them)\n");
        told = true;
    }
    if (const char* e = getenv("SVO_FSHARE")) {
        const int v = atoi(e);
        if (v >= 10 && v <= 100) c->f_share = v;
    }
    if (const char* e = getenv("SVO_TSHARE")) {
        const int v = atoi(e);
        if (v == 0 || (v >= 10 && v <= 100)) c->t_share = v;
    }
    if (!c->spec || p->n_seq != 1) want_chain = want_in = 0;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, p->device));
    bool ok = false;
    // want_in == 0: the ingest shares the chain partition (its stream is created there)
    if (want_chain >= 8 && (want_in >= 8 || want_in == 0) && want_chain + want_in + 16 <= prop.multiProcessorCount) {
#define DRV(name)                                                                                      \
    decltype(&name) p_##name = nullptr;                                                                \
    {                                                                                                  \
        cudaDriverEntryPointQueryResult qr;                                                            \
        if (cudaGetDriverEntryPoint(#name, (void**)&p_##name, cudaEnableDefault, &qr) != cudaSuccess) \
            p_##name = nullptr;                                                                        \
    }
        DRV(cuDeviceGetDevResource) DRV(cuDevSmResourceSplitByCount) DRV(cuDevResourceGenerateDesc)
        DRV(cuGreenCtxCreate) DRV(cuGreenCtxStreamCreate)
#undef DRV
        if (p_cuDeviceGetDevResource && p_cuDevSmResourceSplitByCount && p_cuDevResourceGenerateDesc &&
            p_cuGreenCtxCreate && p_cuGreenCtxStreamCreate) {
            // split the GPU into groups of 8 SMs (the granularity the driver accepts) and assemble the three
            // partitions from whole groups; what is left over joins the KLT partition
            CUdevResource all, grp[32], rem;
            unsigned ngrp = 32;
            CUgreenCtx gc[3] = {nullptr, nullptr, nullptr};
            CUdevResourceDesc dc[3];
            CUstream st[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
            const bool dbg = getenv("SVO_DEBUG") != nullptr;
            int step = 0;
            auto TRY = [&](CUresult x) {
                ++step;
                if (ok && x != CUDA_SUCCESS) {
                    ok = false;
                    if (dbg) fprintf(stderr, "svo: green-context step %d failed with CUresult %d\n", step, (int)x);
                }
                return ok;
            };
            ok = true;
            TRY(p_cuDeviceGetDevResource(p->device, &all, CU_DEV_RESOURCE_TYPE_SM));
            if (ok) TRY(p_cuDevSmResourceSplitByCount(grp, &ngrp, &all, &rem, 0, 8));
            const int n_chain = (want_chain + 7) / 8, n_in = (want_in + 7) / 8;
            ok = ok && (int)ngrp >= n_chain + n_in + 2;
            unsigned sm_cnt[3] = {0, 0, 0};
            if (ok) {
                CUdevResource kl[33];
                int nk = 0;
                for (unsigned i = n_chain + n_in; i < ngrp; ++i) kl[nk++] = grp[i];
                if (rem.sm.smCount > 0) kl[nk++] = rem;
                for (int i = 0; i < n_chain; ++i) sm_cnt[0] += grp[i].sm.smCount;
                for (int i = 0; i < n_in; ++i) sm_cnt[1] += grp[n_chain + i].sm.smCount;
                for (int i = 0; i < nk; ++i) sm_cnt[2] += kl[i].sm.smCount;
                TRY(p_cuDevResourceGenerateDesc(&dc[0], grp, n_chain));
                if (ok && n_in > 0) TRY(p_cuDevResourceGenerateDesc(&dc[1], grp + n_chain, n_in));
                if (ok && p_cuDevResourceGenerateDesc(&dc[2], kl, nk) != CUDA_SUCCESS) {
                    // the remainder may not combine with full groups: leave it unused
                    nk -= rem.sm.smCount > 0 ? 1 : 0;
                    sm_cnt[2] -= rem.sm.smCount;
                    TRY(p_cuDevResourceGenerateDesc(&dc[2], kl, nk));
                }
            }
            if (dbg) fprintf(stderr, "svo: SM partitions chain %u ingest %u klt %u (groups %u, ok %d)\n", sm_cnt[0], sm_cnt[1], sm_cnt[2], ngrp, (int)ok);
            for (int i = 0; i < 3 && ok; ++i)
                if (i != 1 || n_in > 0) TRY(p_cuGreenCtxCreate(&gc[i], dc[i], p->device, CU_GREEN_CTX_DEFAULT_STREAM));
            // the pose chain shares its SMs with the ingest: its blocks are placed first (stream priority; SVO_CHAIN_PRIO=0 for equal)
            const int chain_prio = getenv("SVO_CHAIN_PRIO") ? atoi(getenv("SVO_CHAIN_PRIO")) : -1;
            if (ok) TRY(p_cuGreenCtxStreamCreate(&st[0], gc[0], CU_STREAM_NON_BLOCKING, chain_prio));
            if (ok) TRY(p_cuGreenCtxStreamCreate(&st[1], gc[0], CU_STREAM_NON_BLOCKING, chain_prio));
            if (ok) TRY(p_cuGreenCtxStreamCreate(&st[2], n_in > 0 ? gc[1] : gc[0], CU_STREAM_NON_BLOCKING, 0));
            if (ok) TRY(p_cuGreenCtxStreamCreate(&st[3], gc[2], CU_STREAM_NON_BLOCKING, 0));
            if (ok) TRY(p_cuGreenCtxStreamCreate(&st[4], gc[2], CU_STREAM_NON_BLOCKING, -1));
            if (ok) {
                c->stream = (cudaStream_t)st[0];
                c->stream_fb = (cudaStream_t)st[1];
                c->stream_in = (cudaStream_t)st[2];
                c->stream_k = (cudaStream_t)st[3];
                c->stream_k2 = (cudaStream_t)st[4];
                for (int i = 0; i < 3; ++i) c->green[i] = gc[i];
                c->sm_chain = (int)sm_cnt[0];
                c->sm_in = (int)sm_cnt[1];
                c->sm_k = (int)sm_cnt[2];
            }
            cudaGetLastError();
        } else if (getenv("SVO_DEBUG")) {
            fprintf(stderr, "svo: green-context driver entry points not available\n");
        }
    }
    if (ok) CU(cudaStreamCreateWithFlags(&c->stream_full, cudaStreamNonBlocking));
    if (!ok) {
        c->sm_chain = c->sm_in = c->sm_k = 0;
        CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&c->stream_fb, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&c->stream_in, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&c->stream_k, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&c->stream_k2, cudaStreamNonBlocking));
        c->stream_full = c->stream;
    }
    return SVO_OK;
}

static int ctx_init(svo_ctx* c, const svo_params* p, int cap_override) {
    c->p = *p;
    CU(cudaSetDevice(p->device));
    {
        int rc = create_streams(c, p);
        if (rc != SVO_OK) return rc;
    }
    for (int i = 0; i < SVO_EV_RING; ++i) {
        CU(cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_h2d[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_sel[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_td[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_trk[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_spf[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_ts[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_bm[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_fa[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_fb[i], cudaEventDisableTiming));
    }
    CU(cudaStreamCreateWithFlags(&c->stream_h2d, cudaStreamNonBlocking));
    for (int i = 0; i < SVO_RESULT_RING; ++i) {
        c->ev_done[i] = nullptr;
        CU(cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming));
    }
    c->pg = make_pyr_geom(p->width, p->height, p->max_level, p->win_size, channels_of(*p));
    // look-ahead contexts: room for the batch being tracked and the batch being ingested
    // ring slots: 3 are the minimum (previous, current, being ingested); a single-sequence context takes a fourth so
    // that the ingest -- and with it the FAST-candidate pass -- of frame f + 1 need not wait for select(f - 1)
    // (measured with SVO_TSHARE=100: +5 % resident, +11 % end to end); look-ahead contexts hold two batches
    c->pg.n_slots = p->lookahead > 1 ? 2 * p->lookahead + 3 : (p->n_seq == 1 ? SVO_PYR_SLOTS + 1 : SVO_PYR_SLOTS);
    if (const char* e = getenv("SVO_SLOTS")) {
        const int v = atoi(e);
        if (p->lookahead <= 1 && v >= SVO_PYR_SLOTS && v <= 8) c->pg.n_slots = v;
    }
    c->bg = make_bucket_geom(p->height, p->width, p->buckets_along_height, p->buckets_along_width,
                             p->bucket_start_row, p->age_threshold, p->fast_threshold);
    PipeConfig& pc = c->pc;
    memset(&pc, 0, sizeof(pc));
    pc.width = p->width;
    pc.height = p->height;
    pc.features_threshold = p->features_threshold;
    pc.pre_matching_threshold = p->pre_matching_threshold;
    pc.thr_hi = p->fast_threshold;
    pc.thr_lo = p->fast_threshold / 4;
    pc.circular_threshold = p->circular_threshold;
    pc.reproj_thr_sq = (float)((double)p->ransac_reproj_error * (double)p->ransac_reproj_error);
    pc.max_t_norm = p->max_translation_norm;
    pc.max_r_norm = p->max_rotation_norm;
    pc.ransac_iterations = p->ransac_iterations;
    pc.ransac_confidence = (double)p->ransac_confidence;
    pc.lk = make_lk_config(*p, c->pg.nlevels);
    pc.check_bounds = 1;
    pc.age_increment = 1;

    DevArrays& d = c->d;
    memset(&d, 0, sizeof(d));
    d.n_seq = p->n_seq;
    d.n_hyp = p->ransac_hypotheses;
    int cap = c->bg.nb;   // one feature per bucket at most
    if (cap_override > cap) cap = cap_override;
    cap = (cap + 3) & ~3;   // word-aligned per-sequence status bytes
    d.cap = cap;
    const size_t S = (size_t)p->n_seq, C = S * (size_t)cap;
    CU(dev_alloc(c, &d.pyr, (size_t)c->pg.image_bytes * 2 * c->pg.n_slots * S));
    CU(dev_alloc(c, &d.fastkeys, (size_t)2 * c->pg.n_slots * S * c->bg.nb));
    CU(dev_alloc(c, &d.fast_nA, (size_t)c->pg.n_slots * S));
    CU(dev_alloc(c, &d.trackkeys, S * c->bg.nb));
    CU(dev_alloc(c, &d.tile_cnt, S * 16));
    CU(dev_alloc(c, &d.state, S));
    CU(dev_alloc(c, &d.result, S));
    CU(dev_alloc(c, &d.fs_xy, C));  CU(dev_alloc(c, &d.fs_age, C));  CU(dev_alloc(c, &d.fs_str, C));
    CU(dev_alloc(c, &d.bk_xy, C));  CU(dev_alloc(c, &d.bk_age, C));  CU(dev_alloc(c, &d.bk_str, C));
    CU(dev_alloc(c, &d.lk_L1, C));  CU(dev_alloc(c, &d.lk_R1, C));   CU(dev_alloc(c, &d.lk_R0, C));
    CU(dev_alloc(c, &d.lk_C, C));   CU(dev_alloc(c, &d.lk_status, C));
    CU(dev_alloc(c, &d.m_L0, C));   CU(dev_alloc(c, &d.m_L1, C));    CU(dev_alloc(c, &d.m_R0, C));
    CU(dev_alloc(c, &d.m_R1, C));   CU(dev_alloc(c, &d.m_age, C));   CU(dev_alloc(c, &d.m_str, C));
    CU(dev_alloc(c, &d.X3, C * 3));
    CU(dev_alloc(c, &d.X3u, C * 3));
    CU(dev_alloc(c, &d.in_uv, C));
    CU(dev_alloc(c, &d.hyp_pose, S * d.n_hyp * 12));
    CU(dev_alloc(c, &d.hyp_count, S * d.n_hyp));
    CU(dev_alloc(c, &d.inl_idx, C));
    CU(dev_alloc(c, &d.m_src, C));
    CU(dev_alloc(c, &d.klt_next, S));
    d.nb = c->bg.nb;
    CU(dev_alloc(c, &d.specF_pts, (size_t)2 * S * c->bg.nb * 4));
    CU(dev_alloc(c, &d.specF_st, (size_t)2 * S * c->bg.nb));
    CU(dev_alloc(c, &d.specF_it, (size_t)2 * S * c->bg.nb));
    CU(dev_alloc(c, &d.specT_pts, C * 4));
    CU(dev_alloc(c, &d.specT_st, C));
    CU(dev_alloc(c, &d.specT_it, C));
    CU(dev_alloc(c, &d.cntF, S));
    CU(dev_alloc(c, &d.cntT, S));
    CU(dev_alloc(c, &d.fast_work, (size_t)2 * c->pg.n_slots));
    CU(dev_alloc(c, &d.listF, S * c->bg.nb));
    CU(dev_alloc(c, &d.nF, 2 * S));
    CU(dev_alloc(c, &d.predT, (size_t)2 * S * c->bg.nb));
    CU(dev_alloc(c, &d.todo, C));
    CU(dev_alloc(c, &d.n_todo, S));
    CU(dev_alloc(c, &d.job, S));
    CU(dev_alloc(c, &d.hyp_samples, S * d.n_hyp * 5));
    CU(dev_alloc(c, &d.rng_raw, (size_t)SVO_RNG_TABLE));
    {
        std::vector<unsigned> raw(SVO_RNG_TABLE);
        uint64_t state = 0xFFFFFFFFFFFFFFFFull;   // RNG rng((uint64)-1) in RANSACPointSetRegistrator::run
        for (int i = 0; i < SVO_RNG_TABLE; ++i) raw[i] = cv_rng_next(state);
        CU(cudaMemcpyAsync(d.rng_raw, raw.data(), sizeof(unsigned) * SVO_RNG_TABLE, cudaMemcpyHostToDevice, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    // staging ring for host frames: slot f % 3 (the copy of frame f + 1 runs while frame f is being ingested)
    CU(dev_alloc(c, &c->stage_L, (size_t)c->pg.pitch[0] * p->height * S * c->pg.n_slots, false));
    CU(dev_alloc(c, &c->stage_R, (size_t)c->pg.pitch[0] * p->height * S * c->pg.n_slots, false));
    CU(cudaMallocHost((void**)&c->host_result, sizeof(svo_frame_result) * S * SVO_RESULT_RING));
    // initial state: rotation = I, translation = 0, last_transform = I (vo.h:262-264)
    std::vector<SeqState> init(S);
    memset(init.data(), 0, sizeof(SeqState) * S);
    for (size_t s = 0; s < S; ++s) {
        init[s].R[0] = init[s].R[4] = init[s].R[8] = 1.0;
        init[s].last_T[0] = init[s].last_T[5] = init[s].last_T[10] = init[s].last_T[15] = 1.0;
    }
    CU(cudaMemcpyAsync(d.state, init.data(), sizeof(SeqState) * S, cudaMemcpyHostToDevice, c->stream));
    for (int i = 0; i < 7; ++i) CU(cudaEventCreate(&c->ev[i]));
    {
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, p->device));
        // every resident block slot of the partition the kernel runs on, split between the sequences
        const int per_sm = klt_resident_blocks_per_sm();
        const int sm_chain = c->sm_chain > 0 ? c->sm_chain : prop.multiProcessorCount;
        const int sm_k = c->sm_k > 0 ? c->sm_k : prop.multiProcessorCount;
        int per_seq = (sm_chain * per_sm + p->n_seq - 1) / p->n_seq;
        c->klt_blocks = per_seq < 1 ? 1 : per_seq;
        per_seq = (sm_k * klt_spec_resident_blocks_per_sm() + p->n_seq - 1) / p->n_seq;
        c->klt_blocks_k = per_seq < 1 ? 1 : per_seq;
        per_seq = (prop.multiProcessorCount * per_sm + p->n_seq - 1) / p->n_seq;
        c->klt_blocks_full = per_seq < 1 ? 1 : per_seq;
    }
    CU(cudaStreamSynchronize(c->stream));
    return SVO_OK;
}

static svo_ctx* ctx_create(const svo_params* p, int cap_override, bool pipeline_ctx = true) {
    if (validate_params(p) != SVO_OK) return nullptr;
    if (svo_device_count() <= 0) {
        if (g_err.empty()) g_err = "no CUDA device";
        return nullptr;
    }
    svo_ctx* c = new svo_ctx();
    c->pipeline_ctx = pipeline_ctx;
    c->trace_path = pipeline_ctx ? getenv("SVO_TRACE") : nullptr;
    if (ctx_init(c, p, cap_override) != SVO_OK) {
        std::string keep = g_err;
        svo_destroy(c);
        g_err = keep;
        return nullptr;
    }
    return c;
}

extern "C" svo_ctx* svo_create(const svo_params* p) {
    if (p && (long long)p->buckets_along_height * p->buckets_along_width > 16384) {
        fail(SVO_ERR_INVALID, "the pipeline context supports at most 16384 buckets (reference: 92 x 160 = 14720)");
        return nullptr;
    }
    if (p && p->features_per_bucket != 1) {
        fail(SVO_ERR_INVALID, "the pipeline context is built for FEATURES_PER_BUCKET == 1 (vo.h); use svo_bucket_filter for other values");
        return nullptr;
    }
    return ctx_create(p, 0);
}

static void trace_dump(svo_ctx* c);
extern "C" void svo_destroy(svo_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->p.device);
    for (cudaStream_t st : {c->stream_h2d, c->stream_in, c->stream_k, c->stream_k2, c->stream, c->stream_fb})
        if (st) cudaStreamSynchronize(st);
    trace_dump(c);
    for (void* q : c->allocs) cudaFree(q);
    for (int i = 0; i < SVO_EV_RING; ++i) {
        if (c->ev_in[i]) cudaEventDestroy(c->ev_in[i]);
        if (c->ev_h2d[i]) cudaEventDestroy(c->ev_h2d[i]);
        if (c->ev_sel[i]) cudaEventDestroy(c->ev_sel[i]);
        if (c->ev_td[i]) cudaEventDestroy(c->ev_td[i]);
        if (c->ev_trk[i]) cudaEventDestroy(c->ev_trk[i]);
        if (c->ev_spf[i]) cudaEventDestroy(c->ev_spf[i]);
        if (c->ev_ts[i]) cudaEventDestroy(c->ev_ts[i]);
        if (c->ev_bm[i]) cudaEventDestroy(c->ev_bm[i]);
        if (c->ev_fa[i]) cudaEventDestroy(c->ev_fa[i]);
        if (c->ev_fb[i]) cudaEventDestroy(c->ev_fb[i]);
    }
    if (c->stream_full && c->stream_full != c->stream) {
        cudaStreamSynchronize(c->stream_full);
        cudaStreamDestroy(c->stream_full);
    }
    for (cudaStream_t st : {c->stream_k, c->stream_k2, c->stream_fb, c->stream_h2d})
        if (st) cudaStreamDestroy(st);
    for (int i = 0; i < SVO_RESULT_RING; ++i)
        if (c->ev_done[i]) cudaEventDestroy(c->ev_done[i]);
    if (c->stream_in) cudaStreamDestroy(c->stream_in);
    if (c->host_result) cudaFreeHost(c->host_result);
    if (c->stream) {
        for (int i = 0; i < 7; ++i)
            if (c->ev[i]) cudaEventDestroy(c->ev[i]);
        cudaStreamDestroy(c->stream);
    }
    {
        decltype(&cuGreenCtxDestroy) p_destroy = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (c->green[0] && cudaGetDriverEntryPoint("cuGreenCtxDestroy", (void**)&p_destroy, cudaEnableDefault, &qr) == cudaSuccess &&
            p_destroy)
            for (int i = 0; i < 3; ++i)
                if (c->green[i]) p_destroy((CUgreenCtx)c->green[i]);
    }
    delete c;
}

extern "C" int svo_set_projection(svo_ctx* c, const float Pl[12], const float Pr[12]) {
    if (!c || !Pl || !Pr) return fail(SVO_ERR_INVALID, "NULL argument");
    memcpy(c->pc.Pl, Pl, sizeof(float) * 12);
    memcpy(c->pc.Pr, Pr, sizeof(float) * 12);
    // left_camera_matrix = P_left[:, :3] (vo.cpp:16-25); projectPoints reads fx, fy, cx, cy
    c->pc.K.fx = Pl[0];
    c->pc.K.fy = Pl[5];
    c->pc.K.cx = Pl[2];
    c->pc.K.cy = Pl[6];
    c->have_projection = true;
    return SVO_OK;
}

extern "C" void* svo_stream(svo_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" long long svo_launch_count(svo_ctx* c) { return c ? c->launches : 0; }

extern "C" int svo_enable_timers(svo_ctx* c, int on) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    c->timers = on != 0;
    return SVO_OK;
}
extern "C" int svo_reset_timers(svo_ctx* c) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    for (int i = 0; i < 6; ++i) {
        c->stage_ms[i] = 0;
        c->stage_launches[i] = 0;
    }
    return SVO_OK;
}
extern "C" int svo_get_stage_time(svo_ctx* c, int stage, double* ms_out, long long* launches_out) {
    if (!c || stage < 0 || stage >= 6) return fail(SVO_ERR_INVALID, "bad stage");
    if (ms_out) *ms_out = c->stage_ms[stage];
    if (launches_out) *launches_out = c->stage_launches[stage];
    return SVO_OK;
}

// timeline tracing (debug): brackets a launch with timing events on its stream
struct TraceScope {
    svo_ctx* c;
    cudaStream_t st;
    bool on;
    size_t idx;
    TraceScope(svo_ctx* c_, const char* name, int sid, cudaStream_t st_) : c(c_), st(st_), on(false), idx(0) {
        nvtxRangePushA(name);   // an NVTX range around the launches of every stage
        if (!c->trace_path || c->frames < 30 || c->frames >= 46) return;
        svo_ctx::TraceRec r{name, c->frames, sid, nullptr, nullptr};
        cudaEventCreate(&r.a);
        cudaEventCreate(&r.b);
        cudaEventRecord(r.a, st);
        c->trace.push_back(r);
        idx = c->trace.size() - 1;
        on = true;
    }
    ~TraceScope() {
        if (on) cudaEventRecord(c->trace[idx].b, st);
        nvtxRangePop();
    }
};
static void trace_dump(svo_ctx* c) {
    if (!c->trace_path || c->trace.empty()) return;
    FILE* fp = fopen(c->trace_path, "w");
    if (!fp) return;
    fprintf(fp, "name,frame,stream,start_us,end_us\n");
    for (const auto& r : c->trace) {
        float a = 0.f, b = 0.f;
        cudaEventElapsedTime(&a, c->trace[0].a, r.a);
        cudaEventElapsedTime(&b, c->trace[0].a, r.b);
        fprintf(fp, "%s,%lld,%d,%.1f,%.1f\n", r.name, r.frame, r.stream, a * 1e3, b * 1e3);
    }
    fclose(fp);
    for (auto& r : c->trace) {
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    c->trace.clear();
}

// serial frame: every launch of the frame goes to ONE stream that owns the whole GPU (per-stage timers; a
// partitioned context whose caller is not streaming)
static bool serial_frame(const svo_ctx* c, bool speculate) {
    return c->timers || (c->sm_chain > 0 && !(c->spec && speculate));
}
static cudaStream_t ingest_stream(const svo_ctx* c, bool speculate) {
    return serial_frame(c, speculate) ? c->stream_full : c->stream_in;
}

// One frame = a fixed set of launches on four streams (see the file header).  Dependencies of frame f:
//   ingest(f)    [stream_in]  after select(f - 2): pyramid slot / FAST key buffer f % 3 are free again
//   F(f), T(f)   [stream_k]   speculative circular matching of the candidates of frame f: F needs ingest(f),
//                             T needs the bucketed list of f - 1 with its keep bits (klt(f - 1) done); results are a cache keyed by bucket /
//                             matched index that bucket_merge(f) looks up
//   chain(f)     [stream]     bucket_merge (needs the inliers of f - 1 and the cache) -> klt (only what the cache
//                             missed) -> select -> hypotheses -> inliers / feature set
//   refit(f)     [stream_fb]  LM refit, gates, result struct, result copies
// `left`/`right`: device planes, or nullptr when level 0 of the slot was already written (H2D in place).
// to_host: also copy the result structs into the pinned host ring (slot tracked % SVO_RESULT_RING).
static int enqueue_frame(svo_ctx* c, const uint8_t* left, const uint8_t* right, int stride, long long seq_stride,
                         svo_frame_result* results_dev, bool to_host, bool speculate, bool ingested = false) {
    const long long f = c->frames;
    const bool first = f == 0;
    const int NS = c->pg.n_slots;
    const int cur = (int)(f % NS), prev = (int)((f + NS - 1) % NS);
    const int e = (int)(f & SVO_EV_MASK), e1 = (int)((f + SVO_EV_MASK) & SVO_EV_MASK);   // event slots of frame f and f - 1
    // speculation pays when frames are in flight (its kernels then run beside the previous frame's pose chain);
    // a caller that waits for every result before sending the next frame, and the per-stage timers, get the
    // plain serial chain on one stream that owns the whole GPU
    const bool spec = c->spec && !c->timers && speculate;
    const bool tm = serial_frame(c, speculate);
    cudaStream_t sP = tm ? c->stream_full : c->stream;
    cudaStream_t sI = tm ? sP : c->stream_in, sK = tm ? sP : c->stream_k, sB = tm ? sP : c->stream_fb;
    if (tm && !first) CU(cudaStreamWaitEvent(sP, c->ev_fb[e1], 0));   // the previous frame may have run on the partitions
    if (c->timers) cudaEventRecord(c->ev[0], sP);
    int n_ingest = 0;
    if (ingested) {   // pyramids + FAST keys of this frame came with a look-ahead launch set (svo_submit_frames): ev_in[e] is recorded
        CU(cudaStreamWaitEvent(sP, c->ev_in[e], 0));
    } else {
        {
            TraceScope ts(c, "ingest", 0, sI);
            // (programmatic dependent launch where the set owns its SMs: serial frames and batched contexts)
            n_ingest = launch_ingest(c->d, c->pg, c->bg, c->pc, left, right, stride, seq_stride, cur, sI, tm || c->d.n_seq > 1);
        }
        c->launches += n_ingest;
        if (!tm) {
            CU(cudaEventRecord(c->ev_in[e], sI));
            CU(cudaStreamWaitEvent(sP, c->ev_in[e], 0));
        }
    }
    if (spec && !first) {
        // the two candidate classes run side by side, each on its share (f_share) of the partition's block slots: F(f) as soon
        // as image f is there (typically a frame ahead), T(f) the moment every bucketed feature of f - 1 has its keep bit
        const int half = c->klt_blocks_k > 1 ? (int)((long long)c->klt_blocks_k * c->f_share / 100) : 1;
        cudaStream_t sK2 = c->stream_k2;
        // both passes read pyramid slot / FAST keys of frame f - 1 too: ingest(f - 1) normally precedes ingest(f) on
        // the ingest stream, but not when frame f - 1 was a serial frame on stream_full (first submit after a drain)
        CU(cudaStreamWaitEvent(sK, c->ev_in[e1], 0));
        CU(cudaStreamWaitEvent(sK2, c->ev_in[e1], 0));
        CU(cudaStreamWaitEvent(sK, c->ev_in[e], 0));
        // the FAST-candidate cache is double buffered (f & 1): bucket_merge(f - 2) has to be done reading this half (with
        // three ring slots the ingest's own wait implies it)
        if (f >= 2) CU(cudaStreamWaitEvent(sK, c->ev_bm[(f - 2) & SVO_EV_MASK], 0));
        // ... and T(f - 1), the pass on the critical loop, has to be ready first: both passes are persistent, and an F
        // launch that takes its block slots just before T becomes ready leaves T on the remaining 30 % for its whole
        // duration -- a state that then repeats every frame (resident rate 2900 -> 2170 frames/s, seen in 2 of 11 runs).
        // T's stream has the higher priority, so once both are ready T's blocks are placed first.
        CU(cudaStreamWaitEvent(sK, c->ev_ts[e1], 0));
        {
            TraceScope ts(c, "spec_fast", 1, sK);
            launch_klt_spec_fast(c->d, c->pg, c->pc, c->bg, half, prev, cur, prev, (int)(f & 1), sK);
        }
        CU(cudaEventRecord(c->ev_spf[e], sK));
        CU(cudaStreamWaitEvent(sK2, c->ev_in[e], 0));
        CU(cudaStreamWaitEvent(sK2, c->ev_td[e1], 0));   // the bucketed list of f - 1 with its keep bits
        CU(cudaEventRecord(c->ev_ts[e], sK2));
        {
            TraceScope ts(c, "spec_trk", 2, sK2);
            const int tb = c->t_share > 0 ? (int)((long long)c->klt_blocks_k * c->t_share / 100) : c->klt_blocks_k - half;
            launch_klt_spec_trk(c->d, c->pg, c->pc, c->bg, tb > 0 ? tb : 1, prev, cur, (int)(f & 1), sK2);
        }
        CU(cudaEventRecord(c->ev_trk[e], sK2));
        CU(cudaStreamWaitEvent(sP, c->ev_spf[e], 0));
        CU(cudaStreamWaitEvent(sP, c->ev_trk[e], 0));
        c->launches += 3;
    }
    if (c->timers) cudaEventRecord(c->ev[1], sP);
    if (first) {
        launch_first_frame(c->d, sP);
        c->launches += 1;
        CU(cudaEventRecord(c->ev_sel[e], sP));
        CU(cudaEventRecord(c->ev_td[e], sP));
    } else {
        {
            TraceScope ts(c, "bucket", 3, sP);
            launch_bucket_merge(c->d, c->bg, c->pc, prev, (int)(f & 1), spec ? 1 : 0, sP);   // FAST winners of the PREVIOUS left image
        }
        if (c->timers) cudaEventRecord(c->ev[2], sP);
        if (!tm) CU(cudaEventRecord(c->ev_bm[e], sP));
        if (tm) {
            TraceScope ts(c, "klt", 3, sP);
            launch_klt(c->d, c->pg, c->pc, c->klt_blocks_full, prev, cur, sP);
        } else if (spec) {
            // normally nothing is left to track: a small grid on the chain's own SMs
            TraceScope ts(c, "klt_todo", 3, sP);
            launch_klt(c->d, c->pg, c->pc, c->klt_blocks, prev, cur, sP);
        } else {
            // no cache: all bucketed features, on the KLT partition
            CU(cudaEventRecord(c->ev_trk[e], sP));
            CU(cudaStreamWaitEvent(sK, c->ev_trk[e], 0));
            launch_klt(c->d, c->pg, c->pc, c->klt_blocks_k, prev, cur, sK);
            CU(cudaEventRecord(c->ev_trk[e], sK));
            CU(cudaStreamWaitEvent(sP, c->ev_trk[e], 0));
        }
        if (!tm) CU(cudaEventRecord(c->ev_td[e], sP));   // T(f + 1) may start: it reads the bucketed list + keep bits
        if (c->timers) cudaEventRecord(c->ev[3], sP);
        {
            TraceScope ts(c, "select", 3, sP);
            launch_select(c->d, c->pg, c->pc, prev, cur, prev, (int)(f & 1), sP);
        }
        if (c->timers) cudaEventRecord(c->ev[4], sP);
        CU(cudaEventRecord(c->ev_sel[e], sP));
        {
            TraceScope ts(c, "hyp", 3, sP);
            launch_pnp_hyp(c->d, c->pc, sP);
        }
        if (c->timers) cudaEventRecord(c->ev[5], sP);
        if (!tm) CU(cudaStreamWaitEvent(sP, c->ev_fb[e1], 0));   // the refit of f - 1 still reads the inlier arrays
        {
            TraceScope ts(c, "final_a", 3, sP);
            launch_pnp_final_a(c->d, c->pc, sP);
        }
        c->launches += 8;
    }
    if (!tm) {
        CU(cudaEventRecord(c->ev_fa[e], sP));
        CU(cudaStreamWaitEvent(sB, c->ev_fa[e], 0));
    }
    if (!first) {
        TraceScope ts(c, "final_b", 4, sB);
        launch_pnp_final_b(c->d, c->pc, sB);
    }
    if (c->timers) cudaEventRecord(c->ev[6], sP);
    if (results_dev)
        CU(cudaMemcpyAsync(results_dev, c->d.result, sizeof(svo_frame_result) * c->d.n_seq, cudaMemcpyDeviceToDevice, sB));
    if (to_host) {
        const int r = (int)(c->tracked % SVO_RESULT_RING);
        CU(cudaMemcpyAsync(c->host_result + (size_t)r * c->d.n_seq, c->d.result, sizeof(svo_frame_result) * c->d.n_seq,
                           cudaMemcpyDeviceToHost, sB));
        CU(cudaEventRecord(c->ev_done[r], sB));
        c->tracked += 1;
    }
    if (!tm) {
        CU(cudaEventRecord(c->ev_fb[e], sB));
    } else {   // a later frame may run on the partitions again: it finds every event of this frame signalled
        for (cudaEvent_t ev : {c->ev_in[e], c->ev_td[e], c->ev_sel[e], c->ev_trk[e], c->ev_spf[e], c->ev_ts[e], c->ev_bm[e], c->ev_fa[e], c->ev_fb[e]}) CU(cudaEventRecord(ev, sP));
        // no speculative pass consumed / refreshed the FAST-pruning predictions: drop them
        if (c->spec) CU(cudaMemsetAsync(c->d.predT, 0, sizeof(int) * 2 * c->d.n_seq * c->d.nb, sP));
    }
    CU(cudaGetLastError());
    if (c->timers) {
        CU(cudaStreamSynchronize(sP));
        const int nst = first ? 1 : 6;
        const int nl[6] = {n_ingest, 3, 1, 1, 2, 2};   // bucket = keys + count + merge
        for (int i = 0; i < nst; ++i) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, c->ev[i], c->ev[i + 1]);
            c->stage_ms[i] += ms;
            c->stage_launches[i] += nl[i];
        }
    }
    c->frames += 1;
    return SVO_OK;
}

// the ingest stream may not start writing slot f % NS / key buffer f % NS before its last reader, frame f - NS + 1
// (which tracks from slot (f - NS) % NS), is done with it and has cleared the keys (select)
static int wait_ring(svo_ctx* c, bool speculate) {
    const long long f = c->frames, back = c->pg.n_slots - 1;
    if (f >= back) CU(cudaStreamWaitEvent(ingest_stream(c, speculate), c->ev_sel[(f - back) & SVO_EV_MASK], 0));
    return SVO_OK;
}

extern "C" int svo_process_frame_batch_dev(svo_ctx* c, const uint8_t* left_dev, const uint8_t* right_dev, int stride,
                                           long long seq_stride, svo_frame_result* results_dev) {
    if (!c || !left_dev || !right_dev) return fail(SVO_ERR_INVALID, "NULL argument");
    if (!c->have_projection) return fail(SVO_ERR_STATE, "svo_set_projection has not been called");
    if (stride < c->p.width * c->pg.cn) return fail(SVO_ERR_INVALID, "stride < width * channels");
    CU(cudaSetDevice(c->p.device));
    int rc = wait_ring(c, true);
    if (rc != SVO_OK) return rc;
    return enqueue_frame(c, left_dev, right_dev, stride, seq_stride, results_dev, false, true);
}

extern "C" int svo_submit_frame_batch(svo_ctx* c, const uint8_t* left, const uint8_t* right, int stride,
                                      long long seq_stride) {
    if (!c || !left || !right) return fail(SVO_ERR_INVALID, "NULL argument");
    if (!c->have_projection) return fail(SVO_ERR_STATE, "svo_set_projection has not been called");
    if (stride < c->p.width * c->pg.cn) return fail(SVO_ERR_INVALID, "stride < width * channels");
    if (c->tracked - c->collected >= SVO_RESULT_RING)
        return fail(SVO_ERR_STATE, "result ring full: call svo_collect before submitting more frames");
    CU(cudaSetDevice(c->p.device));
    const bool in_flight = c->tracked > c->collected;   // streaming caller: an earlier frame is still uncollected
    const long long f = c->frames;
    const int S = c->d.n_seq, W = c->p.width * c->pg.cn /* row bytes */, H = c->p.height, P0 = c->pg.pitch[0];
    const long long plane = (long long)P0 * H;
    // host -> staging slot f % 3 on the copy stream: waits only for the ingest of frame f - 3 (the slot's last
    // reader), so the copy of the next frame runs ahead of everything the GPU is still doing for this one
    uint8_t* sl = c->stage_L + (f % c->pg.n_slots) * plane * S;
    uint8_t* sr = c->stage_R + (f % c->pg.n_slots) * plane * S;
    cudaStream_t sH = c->stream_h2d;
    if (f >= c->pg.n_slots) CU(cudaStreamWaitEvent(sH, c->ev_in[(f - c->pg.n_slots) & SVO_EV_MASK], 0));
    if (S == 1 || seq_stride == (long long)stride * H) {
        CU(cudaMemcpy2DAsync(sl, P0, left, stride, W, (size_t)H * S, cudaMemcpyHostToDevice, sH));
        CU(cudaMemcpy2DAsync(sr, P0, right, stride, W, (size_t)H * S, cudaMemcpyHostToDevice, sH));
    } else {
        for (int s = 0; s < S; ++s) {
            CU(cudaMemcpy2DAsync(sl + s * plane, P0, left + s * seq_stride, stride, W, H, cudaMemcpyHostToDevice, sH));
            CU(cudaMemcpy2DAsync(sr + s * plane, P0, right + s * seq_stride, stride, W, H, cudaMemcpyHostToDevice, sH));
        }
    }
    CU(cudaEventRecord(c->ev_h2d[f & SVO_EV_MASK], sH));
    int rc = wait_ring(c, in_flight);
    if (rc != SVO_OK) return rc;
    cudaStream_t sI = ingest_stream(c, in_flight);
    CU(cudaStreamWaitEvent(sI, c->ev_h2d[f & SVO_EV_MASK], 0));
    return enqueue_frame(c, sl, sr, P0, plane, nullptr, true, in_flight);
}

extern "C" int svo_submit_frame(svo_ctx* c, const uint8_t* left, const uint8_t* right, int stride) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    if (c->d.n_seq != 1) return fail(SVO_ERR_INVALID, "svo_submit_frame needs a context with n_seq == 1");
    return svo_submit_frame_batch(c, left, right, stride, 0);
}

extern "C" int svo_submit_frames(svo_ctx* c, int n_frames, const uint8_t* const* lefts, const uint8_t* const* rights,
                                 int stride) {
    if (!c || !lefts || !rights) return fail(SVO_ERR_INVALID, "NULL argument");
    if (c->d.n_seq != 1 || c->pg.cn != 1) return fail(SVO_ERR_INVALID, "svo_submit_frames needs a single-sequence 8UC1 context");
    if (n_frames < 1 || n_frames > (c->p.lookahead > 1 ? c->p.lookahead : 1))
        return fail(SVO_ERR_INVALID, "n_frames must be 1..svo_params.lookahead");
    if (!c->have_projection) return fail(SVO_ERR_STATE, "svo_set_projection has not been called");
    if (stride < c->p.width) return fail(SVO_ERR_INVALID, "stride < width");
    if (c->tracked - c->collected + n_frames > SVO_RESULT_RING)
        return fail(SVO_ERR_STATE, "result ring full: call svo_collect before submitting more frames");
    for (int i = 0; i < n_frames; ++i)
        if (!lefts[i] || !rights[i]) return fail(SVO_ERR_INVALID, "NULL frame pointer");
    if (n_frames == 1) return svo_submit_frame_batch(c, lefts[0], rights[0], stride, 0);
    CU(cudaSetDevice(c->p.device));
    const bool in_flight = c->tracked > c->collected;
    const long long f = c->frames;
    const int NS = c->pg.n_slots, W = c->p.width, H = c->p.height, P0 = c->pg.pitch[0];
    const long long plane = (long long)P0 * H;
    cudaStream_t sH = c->stream_h2d;
    // staging slots (g % NS) are free once the ingest of frame g - NS has read them
    if (f + n_frames - 1 >= NS) CU(cudaStreamWaitEvent(sH, c->ev_in[(f + n_frames - 1 - NS) & SVO_EV_MASK], 0));
    for (int i = 0; i < n_frames; ++i) {
        const long long slot = (f + i) % NS;
        CU(cudaMemcpy2DAsync(c->stage_L + slot * plane, P0, lefts[i], stride, W, H, cudaMemcpyHostToDevice, sH));
        CU(cudaMemcpy2DAsync(c->stage_R + slot * plane, P0, rights[i], stride, W, H, cudaMemcpyHostToDevice, sH));
    }
    CU(cudaEventRecord(c->ev_h2d[f & SVO_EV_MASK], sH));
    cudaStream_t sI = ingest_stream(c, in_flight);
    // ring slots / FAST key buffers of frames f .. f + n - 1 were last used by frames up to f + n - 1 - NS, whose
    // pyramids frame f + n - NS still reads as "previous": its select must be done
    const long long last_reader = f + n_frames - NS;
    if (last_reader >= 0) CU(cudaStreamWaitEvent(sI, c->ev_sel[last_reader & SVO_EV_MASK], 0));
    if (f >= 1) CU(cudaStreamWaitEvent(sI, c->ev_fb[(f - 1) & SVO_EV_MASK], 0));   // (a serial predecessor may have run elsewhere)
    CU(cudaStreamWaitEvent(sI, c->ev_h2d[f & SVO_EV_MASK], 0));
    if (c->timers) cudaEventRecord(c->ev[0], sI);
    int n_launch = 0;
    for (int done = 0; done < n_frames;) {   // consecutive slots; the ring may wrap once
        const int slot0 = (int)((f + done) % NS);
        const int cnt = n_frames - done < NS - slot0 ? n_frames - done : NS - slot0;
        n_launch += launch_ingest_frames(c->d, c->pg, c->bg, c->pc, c->stage_L + slot0 * plane, c->stage_R + slot0 * plane, P0,
                                         plane, slot0, cnt, sI);
        done += cnt;
    }
    c->launches += n_launch;
    for (int i = 0; i < n_frames; ++i) CU(cudaEventRecord(c->ev_in[(f + i) & SVO_EV_MASK], sI));
    if (c->timers) {
        cudaEventRecord(c->ev[1], sI);
        CU(cudaStreamSynchronize(sI));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
        c->stage_ms[0] += ms;
        c->stage_launches[0] += n_launch;
    }
    for (int i = 0; i < n_frames; ++i) {
        const int rc = enqueue_frame(c, nullptr, nullptr, P0, plane, nullptr, true, in_flight || i > 0, true);
        if (rc != SVO_OK) return rc;
    }
    return SVO_OK;
}

extern "C" int svo_pending(svo_ctx* c) { return c ? (int)(c->tracked - c->collected) : SVO_ERR_INVALID; }

extern "C" int svo_collect(svo_ctx* c, svo_frame_result* results) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    if (c->tracked == c->collected) return fail(SVO_ERR_STATE, "no submitted frame is pending");
    CU(cudaSetDevice(c->p.device));
    const int r = (int)(c->collected % SVO_RESULT_RING), S = c->d.n_seq;
    CU(cudaEventSynchronize(c->ev_done[r]));
    const svo_frame_result* hr = c->host_result + (size_t)r * S;
    int ok = 0;
    for (int s = 0; s < S; ++s) ok += hr[s].success ? 1 : 0;
    if (results) memcpy(results, hr, sizeof(svo_frame_result) * S);
    c->last_collected = r;
    c->collected += 1;
    return ok;
}

extern "C" int svo_process_frame_batch(svo_ctx* c, const uint8_t* left, const uint8_t* right, int stride,
                                       long long seq_stride, svo_frame_result* results) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    if (c->tracked != c->collected) return fail(SVO_ERR_STATE, "submitted frames are pending: svo_collect them first");
    int rc = svo_submit_frame_batch(c, left, right, stride, seq_stride);
    if (rc < 0) return rc;
    return svo_collect(c, results);
}

extern "C" int svo_process_frame(svo_ctx* c, const uint8_t* left, const uint8_t* right, int stride, double T_out[16]) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    if (c->d.n_seq != 1) return fail(SVO_ERR_INVALID, "svo_process_frame needs a context with n_seq == 1");
    int rc = svo_process_frame_batch(c, left, right, stride, 0, nullptr);
    if (rc < 0) return rc;
    const svo_frame_result* hr = c->host_result + (size_t)c->last_collected * c->d.n_seq;
    if (T_out) memcpy(T_out, hr->transform, sizeof(double) * 16);
    return hr->success ? 1 : 0;
}

extern "C" int svo_synchronize(svo_ctx* c) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    CU(cudaSetDevice(c->p.device));
    for (cudaStream_t st : {c->stream_in, c->stream_k, c->stream_k2, c->stream, c->stream_fb, c->stream_full}) CU(cudaStreamSynchronize(st));
    return SVO_OK;
}

extern "C" int svo_get_frame_images(svo_ctx* c, int seq, uint8_t* left_out, uint8_t* right_out, int stride) {
    if (!c || !left_out || !right_out) return fail(SVO_ERR_INVALID, "NULL argument");
    if (seq < 0 || seq >= c->d.n_seq) return fail(SVO_ERR_INVALID, "bad sequence index");
    const int row_bytes = c->p.width * c->pg.cn;
    if (stride < row_bytes) return fail(SVO_ERR_INVALID, "stride < width * channels");
    if (c->frames == 0) return fail(SVO_ERR_STATE, "no frame has been processed yet");
    CU(cudaSetDevice(c->p.device));
    int rc = svo_synchronize(c);
    if (rc != SVO_OK) return rc;
    const int slot = (int)((c->frames - 1) % c->pg.n_slots);
    CU(cudaMemcpy2D(left_out, stride, c->d.pyr + c->pg.base(seq, slot, 0) + c->pg.off[0], c->pg.pitch[0], row_bytes,
                    c->p.height, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy2D(right_out, stride, c->d.pyr + c->pg.base(seq, slot, 1) + c->pg.off[0], c->pg.pitch[0], row_bytes,
                    c->p.height, cudaMemcpyDeviceToHost));
    return SVO_OK;
}

extern "C" int svo_pipeline_info(svo_ctx* c, int out[8]) {
    if (!c || !out) return fail(SVO_ERR_INVALID, "NULL argument");
    CU(cudaSetDevice(c->p.device));
    for (cudaStream_t st : {c->stream_in, c->stream_k, c->stream_k2, c->stream, c->stream_fb, c->stream_full}) CU(cudaStreamSynchronize(st));
    int n_todo = 0;
    CU(cudaMemcpy(&n_todo, c->d.n_todo, sizeof(int), cudaMemcpyDeviceToHost));
    out[0] = c->spec ? 1 : 0;
    out[1] = c->sm_chain;
    out[2] = c->sm_in;
    out[3] = c->sm_k;
    out[4] = n_todo;
    out[5] = c->klt_blocks;
    out[6] = c->klt_blocks_k;
    out[7] = 0;
    if (c->spec && c->frames > 1) {   // FAST candidates of the last speculative pass: tracked | skipped << 16
        const long long nb = c->d.nb;
        const int buf = (int)((c->frames - 1) & 1);
        std::vector<int> list(nb);
        std::vector<uint8_t> st(nb);
        int nf[2] = {0, 0};
        CU(cudaMemcpy(&nf[0], c->d.nF, sizeof(int), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(&nf[1], c->d.nF + c->d.n_seq, sizeof(int), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(list.data(), c->d.listF, sizeof(int) * nb, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(st.data(), c->d.specF_st + (long long)buf * c->d.n_seq * nb, nb, cudaMemcpyDeviceToHost));
        int skipped = 0;
        for (int t = 0; t < nf[0] + nf[1] && t < nb; ++t) skipped += (st[list[t < nf[0] ? t : nb - 1 - (t - nf[0])]] & 0x80) ? 1 : 0;
        out[7] = (nf[0] + nf[1] - skipped) | (skipped << 16);
    }
    return SVO_OK;
}

extern "C" int svo_stream_join(svo_ctx* c) {
    if (!c) return fail(SVO_ERR_INVALID, "NULL ctx");
    if (c->frames == 0 || c->timers) return SVO_OK;
    CU(cudaSetDevice(c->p.device));
    const int e = (int)((c->frames - 1) & SVO_EV_MASK);
    CU(cudaStreamWaitEvent(c->stream, c->ev_fb[e], 0));
    CU(cudaStreamWaitEvent(c->stream, c->ev_in[e], 0));
    if (c->spec && c->frames > 1) CU(cudaStreamWaitEvent(c->stream, c->ev_trk[e], 0));
    return SVO_OK;
}

extern "C" int svo_get_result(svo_ctx* c, int seq, svo_frame_result* out) {
    if (!c || !out || seq < 0 || seq >= c->d.n_seq) return fail(SVO_ERR_INVALID, "bad argument");
    CU(cudaSetDevice(c->p.device));
    for (cudaStream_t st : {c->stream_in, c->stream_k, c->stream_k2, c->stream, c->stream_fb, c->stream_full}) CU(cudaStreamSynchronize(st));
    CU(cudaMemcpy(out, c->d.result + seq, sizeof(svo_frame_result), cudaMemcpyDeviceToHost));
    return SVO_OK;
}

extern "C" int svo_get_tap(svo_ctx* c, int seq, int which, int cap, float* f_out, int* i_out, uint8_t* u8_out) {
    if (!c || seq < 0 || seq >= c->d.n_seq || cap < 0) return fail(SVO_ERR_INVALID, "bad argument");
    CU(cudaSetDevice(c->p.device));
    for (cudaStream_t st : {c->stream_in, c->stream_k, c->stream_k2, c->stream, c->stream_fb, c->stream_full}) CU(cudaStreamSynchronize(st));
    SeqState s;
    CU(cudaMemcpy(&s, c->d.state + seq, sizeof(s), cudaMemcpyDeviceToHost));
    const DevArrays& d = c->d;
    const size_t fo = (size_t)seq * d.cap;
    auto get = [&](void* dst, const void* src, size_t bytes) { return cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost); };
    int n = 0;
    if (which == 0 || which == 1) n = s.n_bucketed;
    else if (which == 2) n = s.n_matched;
    else if (which == 3) n = s.n_feat;
    else return fail(SVO_ERR_INVALID, "bad tap id");
    const int m = n < cap ? n : cap;
    if (m == 0) return n;
    if (which == 0 || which == 3) {
        const float2* xy = which == 0 ? d.bk_xy : d.fs_xy;
        const int* age = which == 0 ? d.bk_age : d.fs_age;
        const int* str = which == 0 ? d.bk_str : d.fs_str;
        if (f_out) CU(get(f_out, xy + fo, sizeof(float2) * m));
        if (i_out) {
            CU(get(i_out, age + fo, sizeof(int) * m));
            CU(get(i_out + cap, str + fo, sizeof(int) * m));
        }
    } else if (which == 1) {
        if (f_out) {
            CU(get(f_out + 0 * 2 * (size_t)cap, d.lk_L1 + fo, sizeof(float2) * m));
            CU(get(f_out + 1 * 2 * (size_t)cap, d.lk_R1 + fo, sizeof(float2) * m));
            CU(get(f_out + 2 * 2 * (size_t)cap, d.lk_R0 + fo, sizeof(float2) * m));
            CU(get(f_out + 3 * 2 * (size_t)cap, d.lk_C + fo, sizeof(float2) * m));
        }
        if (u8_out) CU(get(u8_out, d.lk_status + fo, m));
    } else {
        if (f_out) {
            CU(get(f_out + 0 * 2 * (size_t)cap, d.m_L0 + fo, sizeof(float2) * m));
            CU(get(f_out + 1 * 2 * (size_t)cap, d.m_L1 + fo, sizeof(float2) * m));
            CU(get(f_out + 2 * 2 * (size_t)cap, d.m_R0 + fo, sizeof(float2) * m));
            CU(get(f_out + 3 * 2 * (size_t)cap, d.m_R1 + fo, sizeof(float2) * m));
            CU(get(f_out + 4 * 2 * (size_t)cap, d.X3 + fo * 3, sizeof(float) * 3 * m));
        }
        if (i_out) CU(get(i_out, d.inl_idx + fo, sizeof(int) * (s.n_inliers < cap ? s.n_inliers : cap)));
    }
    return n;
}

// ===========================================================================
// Stage-level entry points.  Each call stages its operands in fresh device buffers,
// runs the same kernels the pipeline uses, and copies the result back.
namespace {
struct Scratch {
    std::vector<void*> ptrs;
    cudaStream_t st = nullptr;
    ~Scratch() {
        for (void* p : ptrs) cudaFree(p);
        if (st) cudaStreamDestroy(st);
    }
    template <class T>
    cudaError_t alloc(T** p, size_t count) {
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(q);
        *p = (T*)q;
        return e;
    }
};
int need_device() {
    int n = svo_device_count();
    if (n <= 0) return n < 0 ? n : fail(SVO_ERR_NO_DEVICE, "no CUDA device");
    return SVO_OK;
}
}  // namespace

extern "C" int svo_fast_detect(const uint8_t* img, int width, int height, int stride, int threshold, float* xy_out,
                               float* resp_out, int cap) {
    if (!img || width < 7 || height < 7 || stride < width || threshold < 1 || threshold > 254 || cap < 0)
        return fail(SVO_ERR_INVALID, "bad argument");
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    const int pitch = (width + 127) & ~127;
    uint8_t *dimg, *dmap;
    int *rows, *dn;
    float2* dxy;
    float* dresp;
    CU(s.alloc(&dimg, (size_t)pitch * height));
    CU(s.alloc(&dmap, (size_t)pitch * height));
    CU(s.alloc(&rows, (size_t)height + 1));
    CU(s.alloc(&dn, 1));
    CU(s.alloc(&dxy, (size_t)cap));
    CU(s.alloc(&dresp, (size_t)cap));
    CU(cudaMemcpy2DAsync(dimg, pitch, img, stride, width, height, cudaMemcpyHostToDevice, s.st));
    BucketGeom bg;
    memset(&bg, 0, sizeof(bg));
    launch_fast(dimg, 0, width, height, pitch, 1, threshold, threshold, dmap, 0, pitch, nullptr, nullptr, bg, s.st);
    launch_compact_keypoints(dmap, width, height, pitch, threshold, rows, dxy, dresp, cap, dn, s.st);
    int n = 0;
    CU(cudaMemcpyAsync(&n, dn, sizeof(int), cudaMemcpyDeviceToHost, s.st));
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    const int m = n < cap ? n : cap;
    if (m > 0) {
        if (xy_out) CU(cudaMemcpy(xy_out, dxy, sizeof(float2) * m, cudaMemcpyDeviceToHost));
        if (resp_out) CU(cudaMemcpy(resp_out, dresp, sizeof(float) * m, cudaMemcpyDeviceToHost));
    }
    return n;
}

extern "C" int svo_bucket_filter(float* xy, int* ages, int* strengths, int n, int image_height, int image_width,
                                 int buckets_along_height, int buckets_along_width, int bucket_start_row,
                                 int features_per_bucket, int age_threshold, int fast_threshold) {
    if (n < 0 || (n > 0 && (!xy || !ages || !strengths)) || image_height < 1 || image_width < 1 ||
        buckets_along_height < 1 || buckets_along_width < 1 || features_per_bucket < 0 ||
        (long long)buckets_along_height * buckets_along_width * (features_per_bucket > 0 ? features_per_bucket : 1) > (1ll << 28))
        return fail(SVO_ERR_INVALID, "bad argument");
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    BucketGeom bg = make_bucket_geom(image_height, image_width, buckets_along_height, buckets_along_width,
                                     bucket_start_row, age_threshold, fast_threshold);
    const int pb = features_per_bucket;
    float2 *dxy, *oxy;
    int *dage, *dstr, *oage, *ostr, *scratch, *dn;
    const size_t out_cap = (size_t)n;
    CU(s.alloc(&dxy, (size_t)n));   CU(s.alloc(&dage, (size_t)n));   CU(s.alloc(&dstr, (size_t)n));
    CU(s.alloc(&oxy, out_cap));     CU(s.alloc(&oage, out_cap));     CU(s.alloc(&ostr, out_cap));
    CU(s.alloc(&scratch, (size_t)n + 2 * (size_t)bg.nb + 1 + (size_t)bg.nb * (pb > 0 ? pb : 1)));
    CU(s.alloc(&dn, 1));
    if (n > 0) {
        CU(cudaMemcpyAsync(dxy, xy, sizeof(float2) * n, cudaMemcpyHostToDevice, s.st));
        CU(cudaMemcpyAsync(dage, ages, sizeof(int) * n, cudaMemcpyHostToDevice, s.st));
        CU(cudaMemcpyAsync(dstr, strengths, sizeof(int) * n, cudaMemcpyHostToDevice, s.st));
    }
    launch_bucket_generic(dxy, dage, dstr, n, bg, pb, scratch, oxy, oage, ostr, dn, s.st);
    int m = 0;
    CU(cudaMemcpyAsync(&m, dn, sizeof(int), cudaMemcpyDeviceToHost, s.st));
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    if (m > 0) {
        CU(cudaMemcpy(xy, oxy, sizeof(float2) * m, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(ages, oage, sizeof(int) * m, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(strengths, ostr, sizeof(int) * m, cudaMemcpyDeviceToHost));
    }
    return m;
}

// device pyramid built with the generic pyrdown kernel
namespace {
struct DevPyr {
    PyrGeom pg;
    uint8_t* base = nullptr;
    LkPyramid lk() const {
        LkPyramid p;
        for (int l = 0; l < SVO_MAX_LEVELS; ++l) {
            p.lv[l].img = base + pg.off[l];
            p.lv[l].w = pg.w[l];
            p.lv[l].h = pg.h[l];
            p.lv[l].pitch = pg.pitch[l];
        }
        return p;
    }
};
int build_dev_pyr(Scratch& s, DevPyr& out, const uint8_t* img, int w, int h, int stride, int max_level, int win,
                  int cn = 1) {
    out.pg = make_pyr_geom(w, h, max_level, win, cn);
    CU(s.alloc(&out.base, (size_t)out.pg.image_bytes));
    CU(cudaMemcpy2DAsync(out.base, out.pg.pitch[0], img, stride, (size_t)w * cn, h, cudaMemcpyHostToDevice, s.st));
    for (int l = 1; l < out.pg.nlevels && cn != 1; ++l)
        launch_pyrdown_cn(out.base + out.pg.off[l - 1], out.pg.w[l - 1], out.pg.h[l - 1], out.pg.pitch[l - 1],
                          out.base + out.pg.off[l], out.pg.w[l], out.pg.h[l], out.pg.pitch[l], cn, s.st);
    for (int l = 1; l < out.pg.nlevels && cn == 1; ++l)
        launch_pyrdown(out.base + out.pg.off[l - 1], 0, out.pg.w[l - 1], out.pg.h[l - 1], out.pg.pitch[l - 1],
                       out.base + out.pg.off[l], 0, out.pg.w[l], out.pg.h[l], out.pg.pitch[l], 1, s.st);
    return SVO_OK;
}
}  // namespace

extern "C" int svo_build_pyramid(const uint8_t* img, int width, int height, int stride, int max_level, int win_size,
                                 uint8_t* out, long long out_cap) {
    if (!img || width < 1 || height < 1 || stride < width || max_level < 0 || max_level >= SVO_MAX_LEVELS)
        return fail(SVO_ERR_INVALID, "bad argument");
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    DevPyr p;
    rc = build_dev_pyr(s, p, img, width, height, stride, max_level, win_size);
    if (rc != SVO_OK) return rc;
    long long need = 0;
    for (int l = 1; l < p.pg.nlevels; ++l) need += (long long)p.pg.w[l] * p.pg.h[l];
    if (need > out_cap) return fail(SVO_ERR_CAPACITY, "pyramid output buffer too small");
    long long o = 0;
    for (int l = 1; l < p.pg.nlevels; ++l) {
        CU(cudaMemcpy2DAsync(out + o, p.pg.w[l], p.base + p.pg.off[l], p.pg.pitch[l], p.pg.w[l], p.pg.h[l],
                             cudaMemcpyDeviceToHost, s.st));
        o += (long long)p.pg.w[l] * p.pg.h[l];
    }
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    return p.pg.nlevels;
}

extern "C" int svo_lk_track(const svo_params* p, const uint8_t* prev_img, const uint8_t* next_img, int stride,
                            const float* pts_in, int n, float* pts_out, uint8_t* status_out) {
    int rc = validate_params(p);
    if (rc != SVO_OK) return rc;
    if (!prev_img || !next_img || n < 0 || (n > 0 && (!pts_in || !pts_out || !status_out)) ||
        stride < p->width * channels_of(*p))
        return fail(SVO_ERR_INVALID, "bad argument");
    rc = need_device();
    if (rc != SVO_OK) return rc;
    if (n == 0) return 0;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    DevPyr I, J;
    const int cn = channels_of(*p);
    rc = build_dev_pyr(s, I, prev_img, p->width, p->height, stride, p->max_level, p->win_size, cn);
    if (rc != SVO_OK) return rc;
    rc = build_dev_pyr(s, J, next_img, p->width, p->height, stride, p->max_level, p->win_size, cn);
    if (rc != SVO_OK) return rc;
    float2 *din, *dout;
    uint8_t* dst;
    CU(s.alloc(&din, (size_t)n));
    CU(s.alloc(&dout, (size_t)n));
    CU(s.alloc(&dst, (size_t)n));
    CU(cudaMemcpyAsync(din, pts_in, sizeof(float2) * n, cudaMemcpyHostToDevice, s.st));
    if (cn != 1) launch_lk_single_cn(I.lk(), J.lk(), make_lk_config(*p, I.pg.nlevels), cn, din, n, dout, dst, s.st);
    else launch_lk_single(I.lk(), J.lk(), make_lk_config(*p, I.pg.nlevels), din, n, dout, dst, s.st);
    CU(cudaMemcpyAsync(pts_out, dout, sizeof(float2) * n, cudaMemcpyDeviceToHost, s.st));
    CU(cudaMemcpyAsync(status_out, dst, n, cudaMemcpyDeviceToHost, s.st));
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    int ok = 0;
    for (int i = 0; i < n; ++i) ok += status_out[i] ? 1 : 0;
    return ok;
}

// Stateless circular matching / matchingFeatures on top of a private one-sequence context.
static int run_matching(const svo_params* p, const uint8_t* L0, const uint8_t* R0, const uint8_t* L1,
                        const uint8_t* R1, int stride, bool detect, bool check_bounds, float* fs_xy, int* fs_age,
                        int* fs_str, int* n_fs_io, int fs_cap, const float* pL0_in, int n_in, float* pL0, float* pR0,
                        float* pL1, float* pR1, uint8_t* keep_out) {
    svo_params q = *p;
    q.n_seq = 1;
    q.features_per_bucket = 1;
    // select_kernel / bucket_merge_kernel pack counts into 16-bit halves: the same bounds svo_create enforces
    if (n_in > 32767 || fs_cap > 32767) return fail(SVO_ERR_INVALID, "at most 32767 features per call");
    if ((long long)q.buckets_along_height * q.buckets_along_width > 16384)
        return fail(SVO_ERR_INVALID, "at most 16384 buckets (reference: 92 x 160 = 14720)");
    svo_ctx* c = ctx_create(&q, n_in > fs_cap ? n_in : fs_cap, false);
    if (!c) return SVO_ERR_CUDA;
    struct Guard {
        svo_ctx* c;
        ~Guard() { svo_destroy(c); }
    } guard{c};
    c->pc.check_bounds = check_bounds ? 1 : 0;
    c->pc.age_increment = 0;
    const int W = q.width * c->pg.cn /* row bytes */, H = q.height, P0 = c->pg.pitch[0];
    const long long plane = (long long)P0 * H;
    DevArrays& d = c->d;
    cudaStream_t st = c->stream;
    auto up = [&](const uint8_t* l, const uint8_t* r) -> int {
        CU(cudaMemcpy2DAsync(c->stage_L, P0, l, stride, W, H, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpy2DAsync(c->stage_R, P0, r, stride, W, H, cudaMemcpyHostToDevice, st));
        return SVO_OK;
    };
    int rc = up(L0, R0);
    if (rc != SVO_OK) return rc;
    launch_ingest(d, c->pg, c->bg, c->pc, c->stage_L, c->stage_R, P0, plane, 0, st);
    launch_first_frame(d, st);
    rc = up(L1, R1);
    if (rc != SVO_OK) return rc;
    launch_ingest(d, c->pg, c->bg, c->pc, c->stage_L, c->stage_R, P0, plane, 1, st);
    SeqState s;
    CU(cudaMemcpyAsync(&s, d.state, sizeof(s), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (detect) {
        const int n0 = *n_fs_io;
        if (n0 > 0) {
            CU(cudaMemcpyAsync(d.fs_xy, fs_xy, sizeof(float2) * n0, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(d.fs_age, fs_age, sizeof(int) * n0, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(d.fs_str, fs_str, sizeof(int) * n0, cudaMemcpyHostToDevice, st));
        }
        s.n_feat = n0;
        CU(cudaMemcpyAsync(d.state, &s, sizeof(s), cudaMemcpyHostToDevice, st));
        launch_bucket_merge(d, c->bg, c->pc, 0, 0, 0, st);
    } else {
        if (n_in > 0) CU(cudaMemcpyAsync(d.bk_xy, pL0_in, sizeof(float2) * n_in, cudaMemcpyHostToDevice, st));
        s.n_bucketed = n_in;
        CU(cudaMemcpyAsync(d.state, &s, sizeof(s), cudaMemcpyHostToDevice, st));
        // no bucketing pass here: queue every given point for klt_kernel
        std::vector<int> iota(n_in > 0 ? n_in : 1);
        for (int i = 0; i < n_in; ++i) iota[i] = i;
        CU(cudaMemcpyAsync(d.todo, iota.data(), sizeof(int) * (n_in > 0 ? n_in : 1), cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d.n_todo, &n_in, sizeof(int), cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
    }
    CU(cudaMemsetAsync(d.klt_next, 0, sizeof(int), st));
    launch_klt(d, c->pg, c->pc, c->klt_blocks, 0, 1, st);
    launch_select(d, c->pg, c->pc, 0, 1, detect ? 0 : -1, -1, st);
    CU(cudaMemcpyAsync(&s, d.state, sizeof(s), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    const int m = s.n_matched;
    if (m > 0) {
        if (pL0) CU(cudaMemcpy(pL0, d.m_L0, sizeof(float2) * m, cudaMemcpyDeviceToHost));
        if (pR0) CU(cudaMemcpy(pR0, d.m_R0, sizeof(float2) * m, cudaMemcpyDeviceToHost));
        if (pL1) CU(cudaMemcpy(pL1, d.m_L1, sizeof(float2) * m, cudaMemcpyDeviceToHost));
        if (pR1) CU(cudaMemcpy(pR1, d.m_R1, sizeof(float2) * m, cudaMemcpyDeviceToHost));
        if (detect && m <= fs_cap) {
            CU(cudaMemcpy(fs_xy, d.fs_xy, sizeof(float2) * m, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(fs_age, d.fs_age, sizeof(int) * m, cudaMemcpyDeviceToHost));
            CU(cudaMemcpy(fs_str, d.fs_str, sizeof(int) * m, cudaMemcpyDeviceToHost));
        }
    }
    if (keep_out && s.n_bucketed > 0) {
        std::vector<uint8_t> stt(s.n_bucketed);
        CU(cudaMemcpy(stt.data(), d.lk_status, s.n_bucketed, cudaMemcpyDeviceToHost));
        for (int i = 0; i < s.n_bucketed; ++i) keep_out[i] = (stt[i] & 16) ? 1 : 0;
    }
    if (detect) {
        if (m > fs_cap) return fail(SVO_ERR_CAPACITY, "feature-set buffer too small");
        *n_fs_io = m;
    }
    return m;
}

extern "C" int svo_circular_match(const svo_params* p, const uint8_t* L0, const uint8_t* R0, const uint8_t* L1,
                                  const uint8_t* R1, int stride, float* pL0, int n, float* pR0, float* pL1, float* pR1,
                                  uint8_t* keep_out) {
    int rc = validate_params(p);
    if (rc != SVO_OK) return rc;
    if (!L0 || !R0 || !L1 || !R1 || n < 0 || (n > 0 && !pL0) || stride < p->width * channels_of(*p))
        return fail(SVO_ERR_INVALID, "bad argument");
    rc = need_device();
    if (rc != SVO_OK) return rc;
    if (n == 0) return 0;   // vo.cpp:179-181
    return run_matching(p, L0, R0, L1, R1, stride, false, false, nullptr, nullptr, nullptr, nullptr, 0, pL0, n, pL0,
                        pR0, pL1, pR1, keep_out);
}

extern "C" int svo_matching_features(const svo_params* p, const uint8_t* L0, const uint8_t* R0, const uint8_t* L1,
                                     const uint8_t* R1, int stride, float* fs_xy, int* fs_age, int* fs_str, int* n_fs,
                                     int fs_cap, float* pL0, float* pR0, float* pL1, float* pR1) {
    int rc = validate_params(p);
    if (rc != SVO_OK) return rc;
    if (!L0 || !R0 || !L1 || !R1 || !n_fs || *n_fs < 0 || fs_cap < *n_fs || !fs_xy || !fs_age || !fs_str ||
        stride < p->width * channels_of(*p))
        return fail(SVO_ERR_INVALID, "bad argument");
    rc = need_device();
    if (rc != SVO_OK) return rc;
    return run_matching(p, L0, R0, L1, R1, stride, true, true, fs_xy, fs_age, fs_str, n_fs, fs_cap, nullptr, 0, pL0,
                        pR0, pL1, pR1, nullptr);
}

extern "C" int svo_find_close_points(const float* p1, const float* p2, int n, float threshold, uint8_t* ok_out) {
    if (n < 0 || (n > 0 && (!p1 || !p2 || !ok_out))) return fail(SVO_ERR_INVALID, "bad argument");
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    if (n == 0) return 0;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    float2 *a, *b;
    uint8_t* ok;
    CU(s.alloc(&a, (size_t)n));
    CU(s.alloc(&b, (size_t)n));
    CU(s.alloc(&ok, (size_t)n));
    CU(cudaMemcpyAsync(a, p1, sizeof(float2) * n, cudaMemcpyHostToDevice, s.st));
    CU(cudaMemcpyAsync(b, p2, sizeof(float2) * n, cudaMemcpyHostToDevice, s.st));
    launch_close_points(a, b, n, threshold, ok, s.st);
    CU(cudaMemcpyAsync(ok_out, ok, n, cudaMemcpyDeviceToHost, s.st));
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    int k = 0;
    for (int i = 0; i < n; ++i) k += ok_out[i] ? 1 : 0;
    return k;
}

extern "C" int svo_compact(void* data, int elem_bytes, int n, const uint8_t* keep) {
    if (n < 0 || elem_bytes < 1 || (n > 0 && (!data || !keep))) return fail(SVO_ERR_INVALID, "bad argument");
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    if (n == 0) return 0;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    uint8_t *src, *dst, *k;
    int* dn;
    CU(s.alloc(&src, (size_t)n * elem_bytes));
    CU(s.alloc(&dst, (size_t)n * elem_bytes));
    CU(s.alloc(&k, (size_t)n));
    CU(s.alloc(&dn, 1));
    CU(cudaMemcpyAsync(src, data, (size_t)n * elem_bytes, cudaMemcpyHostToDevice, s.st));
    CU(cudaMemcpyAsync(k, keep, n, cudaMemcpyHostToDevice, s.st));
    launch_compact_bytes(src, dst, elem_bytes, n, k, dn, s.st);
    int m = 0;
    CU(cudaMemcpyAsync(&m, dn, sizeof(int), cudaMemcpyDeviceToHost, s.st));
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    if (m > 0) CU(cudaMemcpy(data, dst, (size_t)m * elem_bytes, cudaMemcpyDeviceToHost));
    return m;
}

extern "C" int svo_triangulate(const float Pl[12], const float Pr[12], const float* pL, const float* pR, int n,
                               float* X3_out) {
    if (!Pl || !Pr || n < 0 || (n > 0 && (!pL || !pR || !X3_out))) return fail(SVO_ERR_INVALID, "bad argument");
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    if (n == 0) return 0;
    Scratch s;
    CU(cudaStreamCreate(&s.st));
    float *dPl, *dPr, *dX;
    float2 *a, *b;
    CU(s.alloc(&dPl, 12));
    CU(s.alloc(&dPr, 12));
    CU(s.alloc(&a, (size_t)n));
    CU(s.alloc(&b, (size_t)n));
    CU(s.alloc(&dX, (size_t)n * 3));
    CU(cudaMemcpyAsync(dPl, Pl, sizeof(float) * 12, cudaMemcpyHostToDevice, s.st));
    CU(cudaMemcpyAsync(dPr, Pr, sizeof(float) * 12, cudaMemcpyHostToDevice, s.st));
    CU(cudaMemcpyAsync(a, pL, sizeof(float2) * n, cudaMemcpyHostToDevice, s.st));
    CU(cudaMemcpyAsync(b, pR, sizeof(float2) * n, cudaMemcpyHostToDevice, s.st));
    launch_triangulate(dPl, dPr, a, b, n, dX, s.st);
    CU(cudaMemcpyAsync(X3_out, dX, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, s.st));
    CU(cudaStreamSynchronize(s.st));
    CU(cudaGetLastError());
    return n;
}

extern "C" int svo_camera_to_world(const svo_params* p, const float K[9], const float* cam_pts, const float* world_pts,
                                   int n, double R[9], double t[3], int* inliers_out, int* success_out) {
    if (!p || !K || !R || !t || n < 4 || !cam_pts || !world_pts) return fail(SVO_ERR_INVALID, "bad argument (n >= 4 required)");
    svo_params q = *p;
    q.n_seq = 1;
    q.features_per_bucket = 1;
    q.features_threshold = 0;   // the >15 gate belongs to stereo_callback, not to cameraToWorld
    if (q.width < 16) q.width = 64;
    if (q.height < 16) q.height = 64;
    int rc = validate_params(&q);
    if (rc != SVO_OK) return rc;
    rc = need_device();
    if (rc != SVO_OK) return rc;
    svo_ctx* c = ctx_create(&q, n, false);
    if (!c) return SVO_ERR_CUDA;
    struct Guard {
        svo_ctx* c;
        ~Guard() { svo_destroy(c); }
    } guard{c};
    c->pc.K.fx = K[0];
    c->pc.K.fy = K[4];
    c->pc.K.cx = K[2];
    c->pc.K.cy = K[5];
    c->pc.max_t_norm = INFINITY;   // the motion gates belong to stereo_callback
    c->pc.max_r_norm = INFINITY;
    DevArrays& d = c->d;
    cudaStream_t st = c->stream;
    SeqState s;
    CU(cudaMemcpy(&s, d.state, sizeof(s), cudaMemcpyDeviceToHost));
    memcpy(s.R, R, sizeof(double) * 9);
    memcpy(s.t, t, sizeof(double) * 3);
    s.n_matched = n;
    s.n_feat = n;
    s.frame_id = 1;
    CU(cudaMemcpyAsync(d.state, &s, sizeof(s), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d.m_L1, cam_pts, sizeof(float2) * n, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d.X3, world_pts, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st));
    launch_pnp_samples(d, st);
    launch_pnp(d, c->pc, st);
    CU(cudaMemcpyAsync(&s, d.state, sizeof(s), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    if (success_out) *success_out = s.pnp_ok;
    if (s.pnp_ok) {
        memcpy(R, s.R, sizeof(double) * 9);
        memcpy(t, s.t, sizeof(double) * 3);
        if (inliers_out && s.n_inliers > 0)
            CU(cudaMemcpy(inliers_out, d.inl_idx, sizeof(int) * s.n_inliers, cudaMemcpyDeviceToHost));
        return s.n_inliers;
    }
    return 0;
}

extern "C" int svo_get_inverse_transform(const double R[9], const double t[3], double T_out[16]) {
    if (!R || !t || !T_out) return fail(SVO_ERR_INVALID, "NULL argument");
    // general 4x4 inverse of [R t; 0 0 0 1] by Gauss-Jordan with partial pivoting (cv::Mat::inv, DECOMP_LU)
    double A[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 8; ++j) A[i][j] = 0.0;
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) A[i][j] = R[i * 3 + j];
        A[i][3] = t[i];
    }
    A[3][3] = 1.0;
    for (int i = 0; i < 4; ++i) A[i][4 + i] = 1.0;
    for (int col = 0; col < 4; ++col) {
        int piv = col;
        for (int r = col + 1; r < 4; ++r)
            if (fabs(A[r][col]) > fabs(A[piv][col])) piv = r;
        if (fabs(A[piv][col]) < 1e-300) {
            for (int i = 0; i < 16; ++i) T_out[i] = 0.0;   // cv::Mat::inv of a singular matrix returns zeros
            return SVO_OK;
        }
        if (piv != col)
            for (int j = 0; j < 8; ++j) {
                double tmp = A[col][j];
                A[col][j] = A[piv][j];
                A[piv][j] = tmp;
            }
        const double inv = 1.0 / A[col][col];
        for (int j = 0; j < 8; ++j) A[col][j] *= inv;
        for (int r = 0; r < 4; ++r)
            if (r != col) {
                const double f = A[r][col];
                if (f != 0.0)
                    for (int j = 0; j < 8; ++j) A[r][j] -= f * A[col][j];
            }
    }
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) T_out[i * 4 + j] = A[i][4 + j];
    return SVO_OK;
}

// ===========================================================================
// svo_fast_detect_noalloc: svo_fast_detect for callers that may not allocate (CUDA graph capture, serving loops).
// No counterpart in the reference.  The response map and the row
// counts of the ordered compaction -- what svo_fast_detect cudaMallocs per call -- live in ONE file-scope device
// array that every GPU loading the library holds until exit; every call on a device shares it.
static __device__ __align__(256) unsigned char g_noalloc_scratch[SVO_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kNoallocMaxDev = 64;
// g_noalloc_scratch is a host-side stand-in here: the device address is looked up once per device (that device
// current) and is what the kernels get.
int noalloc_scratch_ptr(unsigned char** out) {
    static unsigned char* ptrs[kNoallocMaxDev] = {};
    static std::mutex mu;
    int dev = -1;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoallocMaxDev) return fail(SVO_ERR_INVALID, "svo_fast_detect_noalloc: device ordinal beyond 63");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(SVO_ERR_CUDA, "svo_fast_detect_noalloc: cudaGetSymbolAddress returned a null address");
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    *out = ptrs[dev];
    return SVO_OK;
}
}  // namespace

extern "C" int svo_fast_detect_noalloc(const uint8_t* img_dev, int width, int height, int pitch, int threshold,
                                       float* xy_out_dev, float* resp_out_dev, int cap, int* n_out_dev, void* stream) {
    if (!img_dev || !xy_out_dev || !resp_out_dev || !n_out_dev || width < 7 || height < 7 || pitch < width ||
        threshold < 1 || threshold > 254 || cap < 0)
        return fail(SVO_ERR_INVALID, "bad argument");
    const long long map_pitch = ((long long)width + 127) & ~127ll;
    const long long map_bytes = (map_pitch * height + 255) & ~255ll;            // response map, then the row counts
    const long long need = map_bytes + (long long)sizeof(int) * (height + 1);
    if (need > (long long)SVO_NOALLOC_SCRATCH_BYTES) {
        char msg[192];
        snprintf(msg, sizeof(msg),
                 "svo_fast_detect_noalloc: a %d x %d image needs %lld bytes of scratch, the static array holds %lld "
                 "(SVO_NOALLOC_SCRATCH_BYTES); use svo_fast_detect",
                 width, height, need, (long long)SVO_NOALLOC_SCRATCH_BYTES);
        return fail(SVO_ERR_CAPACITY, msg);
    }
    int rc = need_device();
    if (rc != SVO_OK) return rc;
    unsigned char* scratch = nullptr;
    rc = noalloc_scratch_ptr(&scratch);
    if (rc != SVO_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint8_t* dmap = scratch;
    int* rows = reinterpret_cast<int*>(scratch + map_bytes);
    // Nothing of the last call survives into this one's result: fast_kernel writes every pixel of the width x height
    // map, kp_rowcount_kernel every row count, scan_kernel the total -- so there is nothing to clear.
    BucketGeom bg;
    memset(&bg, 0, sizeof(bg));
    launch_fast(img_dev, 0, width, height, pitch, 1, threshold, threshold, dmap, 0, (int)map_pitch, nullptr, nullptr, bg, st);
    launch_compact_keypoints(dmap, width, height, (int)map_pitch, threshold, rows, reinterpret_cast<float2*>(xy_out_dev),
                             resp_out_dev, cap, n_out_dev, st);
    CU(cudaPeekAtLastError());
    return SVO_OK;
}
