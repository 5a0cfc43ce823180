// A7/A8/A10: feature-to-prototype L2 distance, soft weights, label rectification
// (self_training.py:160-165, :139-142, :106-112).
//
// The reference loops over classes and re-reads the whole [N,K,H,W] feature tensor C times; here feat
// is read ONCE.  The contraction is [px, K] x [K, C <= 20]: with so few classes it is bound by HBM and
// the fp32 FMA pipe, not a tensor-core shape (BASELINE.json north_star), so it runs on CUDA cores in the
// reference's direct-difference form (p - f)^2 with packed fp32x2 math (FADD2/FFMA2: two channels per
// issue slot), prototypes broadcast from shared memory as 128-bit loads.
//
// CTA = 8 warps, tile = 64 pixels (lane l owns pixels l and l+32).  The K channels are split across the
// 8 warps (intra-CTA split-K) so that even a single 129x257 image yields > 4000 warps of loads in flight;
// partial squared distances are combined through shared memory, then 64 threads finish sqrt / min /
// softmax / rectify for their pixel.
#include "common.cuh"

namespace tss {

constexpr int kDThreads = 256;
constexpr int kDWarps = 8;
constexpr int kDTile = 64;

struct DistArgs {
    const float* feat;
    const float* protos;
    float* dist_out;
    float* weight_out;
    const float* soft;
    const void* hard;
    int64_t* label_out;
    int64_t N, K, P;
    int C, KC;
    float temperature, train_thred;
    int hard_dtype, ignore;
};

// per-pixel epilogue: weights (A8) and rectified label (A10) from the C distances of pixel p of image n
template <int CT>
__device__ __forceinline__ void distance_epilogue(const DistArgs& a, const float* dsm, int px, int64_t n, int64_t p) {
    float d[CT];
    float dmin = INFINITY;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) {
            d[c] = dsm[c * kDTile + px];
            dmin = fminf(dmin, d[c]);
        }
    const int64_t obase = n * a.C * a.P + p;
    if (a.dist_out) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < a.C) a.dist_out[obase + c * a.P] = d[c];
    }
    if (a.weight_out == nullptr && a.soft == nullptr) return;
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) {
            d[c] = expf(-(d[c] - dmin) * a.temperature);  // softmax(-(d - dmin) * tau); its row maximum is exactly 0
            s += d[c];
        }
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) d[c] = __fdiv_rn(d[c], s);
    if (a.weight_out) {
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < a.C) a.weight_out[obase + c * a.P] = d[c];
    }
    if (a.soft == nullptr || a.label_out == nullptr) return;
    float sf[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) sf[c] = __ldg(a.soft + obase + static_cast<int64_t>(min(c, a.C - 1)) * a.P);  // all in flight
    float best = 0.f, rsum = 0.f;
    int lab = 0;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) {
            const float r = d[c] * sf[c];
            rsum += r;
            if (c == 0 || r > best) { best = r; lab = c; }
        }
    long long out = lab;
    if (__fdiv_rn(best, rsum) < a.train_thred) out = a.ignore;
    if (a.hard != nullptr) {
        long long hl;
        if (a.hard_dtype == TSS_U8) hl = load_label<uint8_t>(a.hard, n * a.P + p);
        else if (a.hard_dtype == TSS_I32) hl = load_label<int32_t>(a.hard, n * a.P + p);
        else hl = load_label<int64_t>(a.hard, n * a.P + p);
        if (hl != a.ignore) out = hl;
    }
    a.label_out[n * a.P + p] = out;
}

// Persistent CTAs walk 64-pixel tiles.  When the whole prototype matrix fits shared memory (KC == K: 156 KB at
// C = 19, K = 2048) it is staged ONCE per CTA; otherwise K is processed in chunks of KC channels that are
// re-staged per tile.
template <int CT>
__global__ void __launch_bounds__(kDThreads, 1) distance_kernel(const DistArgs a, int64_t tiles) {
    extern __shared__ __align__(16) float sm[];
    const int KC = a.KC, PS = KC + 4;  // prototype row stride in shared memory (16-byte aligned rows)
    float* ps = sm;                    // [CT][PS]
    float* red = sm + CT * PS;         // [8][CT][64] split-K partials
    float* dsm = red + kDWarps * CT * kDTile;  // [CT][64] distances
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t tiles_per_img = (a.P + kDTile - 1) / kDTile;
    const bool resident = KC == a.K;
    const int KW = KC / kDWarps;  // channels per warp per chunk (multiple of 4)
    const int kb = warp * KW;

    auto stage = [&](int64_t kc0) {
        for (int i = tid; i < CT * KC; i += kDThreads) {
            const int c = i / KC, kk = i - c * KC;
            ps[c * PS + kk] = (c < a.C) ? a.protos[c * a.K + kc0 + kk] : 0.f;
        }
    };
    if (resident) stage(0);
    __syncthreads();

    for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int64_t n = tile / tiles_per_img;
        const int64_t p0 = (tile - n * tiles_per_img) * kDTile;
        const int64_t pa = min(p0 + lane, a.P - 1), pb = min(p0 + 32 + lane, a.P - 1);
        const float* fimg = a.feat + n * a.K * a.P;
        f32x2 accA[CT], accB[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) { accA[c] = 0ull; accB[c] = 0ull; }

        for (int64_t kc0 = 0; kc0 < a.K; kc0 += KC) {
            if (!resident) {
                __syncthreads();
                stage(kc0);
                __syncthreads();
            }
            const float* fa = fimg + (kc0 + kb) * a.P + pa;
            const float* fb = fimg + (kc0 + kb) * a.P + pb;
            // four channels x two pixels of features against every class: 1 broadcast LDS.128 + 8 packed FP32 ops
            auto compute4 = [&](const float (&r)[8], int kk) {
                const f32x2 A01 = pack2(r[0], r[1]), A23 = pack2(r[2], r[3]), B01 = pack2(r[4], r[5]), B23 = pack2(r[6], r[7]);
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                    const ulonglong2 p = *reinterpret_cast<const ulonglong2*>(ps + c * PS + kb + kk);
                    f32x2 d;
                    d = sub2(p.x, A01); accA[c] = fma2(d, d, accA[c]);
                    d = sub2(p.y, A23); accA[c] = fma2(d, d, accA[c]);
                    d = sub2(p.x, B01); accB[c] = fma2(d, d, accB[c]);
                    d = sub2(p.y, B23); accB[c] = fma2(d, d, accB[c]);
                }
            };
            auto load4 = [&](float (&r)[8], int kk) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    r[j] = ldg_stream_f(fa + (kk + j) * a.P);
                    r[4 + j] = ldg_stream_f(fb + (kk + j) * a.P);
                }
            };
            // register pipeline, three groups deep: the loads of group g+2 are issued before group g is consumed
            // (the kernel is FP32-pipe bound; what has to be hidden is the ~1 us HBM latency with 8 warps per SM)
            float r0[8], r1[8], r2[8];
            load4(r0, 0);
            if (4 < KW) load4(r1, 4);
            int kk = 0;
            for (; kk + 12 <= KW; kk += 12) {
                if (kk + 8 < KW) load4(r2, kk + 8);
                compute4(r0, kk);
                if (kk + 12 < KW) load4(r0, kk + 12);
                compute4(r1, kk + 4);
                if (kk + 16 < KW) load4(r1, kk + 16);
                compute4(r2, kk + 8);
            }
            // tail: 0, 4 or 8 channels left; r0 / r1 already hold them
            if (kk < KW) compute4(r0, kk);
            if (kk + 4 < KW) compute4(r1, kk + 4);
        }
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            float lo, hi;
            unpack2(accA[c], lo, hi);
            red[(warp * CT + c) * kDTile + lane] = lo + hi;
            unpack2(accB[c], lo, hi);
            red[(warp * CT + c) * kDTile + 32 + lane] = lo + hi;
        }
        __syncthreads();
        {
            const int px = tid & (kDTile - 1);
            for (int c = tid >> 6; c < a.C; c += kDThreads / kDTile) {
                float s = 0.f;
#pragma unroll
                for (int w = 0; w < kDWarps; ++w) s += red[(w * CT + c) * kDTile + px];
                dsm[c * kDTile + px] = sqrtf(s);
            }
        }
        __syncthreads();
        if (tid < kDTile && p0 + tid < a.P) distance_epilogue<CT>(a, dsm, tid, n, p0 + tid);
        // the next tile's writes to red[] / dsm[] come after its own K loop and the __syncthreads above
        // red[] is rewritten only after every thread has passed the second barrier; dsm[] readers (tid < 64) reach
        // the first barrier of the next tile before anyone writes dsm[] again
    }
}

template <int CT>
static int launch_distance(DistArgs& a, cudaStream_t st) {
    // whole prototype matrix resident when it fits; else the largest K chunk 32*g (g | K/32, g <= 16)
    const size_t fixed = sizeof(float) * (static_cast<size_t>(kDWarps) * CT * kDTile + CT * kDTile);
    const size_t budget = 220 * 1024;
    if (fixed + sizeof(float) * CT * (a.K + 4) <= budget) {
        a.KC = static_cast<int>(a.K);
    } else {
        const int64_t m = a.K / 32;
        int g = 1;
        for (int cand = 16; cand >= 1; --cand)
            if (m % cand == 0) { g = cand; break; }
        a.KC = 32 * g;
    }
    const size_t smem = fixed + sizeof(float) * CT * (a.KC + 4);
    cudaError_t e = cudaFuncSetAttribute(distance_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return cuda_status(e, "tss_proto_distance: smem attribute");
    const int64_t tiles = a.N * ((a.P + kDTile - 1) / kDTile);
    int cps = static_cast<int>(budget / (smem + 1024));   // CTAs per SM by shared memory (registers allow 1 x 255 or 2 x 128 ...)
    if (cps > 2) cps = 2;
    if (cps < 1) cps = 1;
    const int64_t slots = static_cast<int64_t>(kNumSMs) * cps;
    const unsigned grid = static_cast<unsigned>(tiles < slots ? tiles : slots);
    distance_kernel<CT><<<grid, kDThreads, smem, st>>>(a, tiles);
    TSS_LAUNCH_CHECK("tss_proto_distance");
    return TSS_OK;
}

}  // namespace tss

extern "C" int tss_proto_distance(const float* feat, const float* protos, int64_t n_img, int64_t channels,
                                  int64_t plane, int num_classes, float temperature, float* dist_out,
                                  float* weight_out, const float* soft, const void* hard, int hard_dtype,
                                  float train_thred, int ignore_index, int64_t* label_out, void* stream) {
    using namespace tss;
    TSS_REQUIRE(feat && protos, TSS_E_ARG, "tss_proto_distance: null pointer");
    TSS_REQUIRE(n_img >= 0 && plane > 0 && channels > 0, TSS_E_ARG, "tss_proto_distance: bad extents");
    TSS_REQUIRE(channels % 32 == 0, TSS_E_ARG, "tss_proto_distance: channels (%lld) must be a multiple of 32", static_cast<long long>(channels));
    TSS_REQUIRE(num_classes >= 1 && num_classes <= 20, TSS_E_CLASSES, "tss_proto_distance: num_classes %d outside [1, 20]", num_classes);
    TSS_REQUIRE(dist_out || weight_out || (soft && label_out), TSS_E_ARG, "tss_proto_distance: no output requested");
    TSS_REQUIRE(hard == nullptr || hard_dtype == TSS_U8 || hard_dtype == TSS_I32 || hard_dtype == TSS_I64, TSS_E_ARG, "tss_proto_distance: bad hard dtype");
    if (n_img == 0) return TSS_OK;
    DistArgs a;
    a.feat = feat; a.protos = protos; a.dist_out = dist_out; a.weight_out = weight_out; a.soft = soft;
    a.hard = hard; a.label_out = label_out; a.N = n_img; a.K = channels; a.P = plane; a.C = num_classes;
    a.temperature = temperature; a.train_thred = train_thred; a.hard_dtype = hard_dtype; a.ignore = ignore_index;
    a.KC = 0;  // chosen per class-count specialisation in launch_distance
    cudaStream_t st = as_stream(stream);
    if (num_classes <= 4) return launch_distance<4>(a, st);
    if (num_classes <= 8) return launch_distance<8>(a, st);
    if (num_classes <= 13) return launch_distance<13>(a, st);
    if (num_classes <= 16) return launch_distance<16>(a, st);
    if (num_classes <= 19) return launch_distance<19>(a, st);
    return launch_distance<20>(a, st);
}
