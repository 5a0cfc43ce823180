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

template <int CT>
__global__ void __launch_bounds__(kDThreads, 2) distance_kernel(const DistArgs a) {
    extern __shared__ __align__(16) float sm[];
    const int KC = a.KC, PS = KC + 4;  // prototype row stride in shared memory (16-byte aligned rows)
    float* ps = sm;                    // [CT][PS]; later reused as red[8][CT][64]
    const int main_floats = max(CT * PS, kDWarps * CT * kDTile);
    float* dsm = sm + main_floats;     // [CT][64] distances
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t tiles_per_img = (a.P + kDTile - 1) / kDTile;
    const int64_t n = blockIdx.x / tiles_per_img;
    const int64_t p0 = (blockIdx.x - n * tiles_per_img) * kDTile;
    const int64_t pa = min(p0 + lane, a.P - 1), pb = min(p0 + 32 + lane, a.P - 1);
    const float* fimg = a.feat + n * a.K * a.P;

    f32x2 accA[CT], accB[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) { accA[c] = 0ull; accB[c] = 0ull; }

    const int KW = KC / kDWarps;  // channels per warp per chunk (multiple of 4)
    for (int64_t kc0 = 0; kc0 < a.K; kc0 += KC) {
        __syncthreads();
        for (int i = tid; i < CT * KC; i += kDThreads) {
            const int c = i / KC, kk = i - c * KC;
            ps[c * PS + kk] = (c < a.C) ? a.protos[c * a.K + kc0 + kk] : 0.f;
        }
        __syncthreads();
        const int kb = warp * KW;
        const float* fa = fimg + (kc0 + kb) * a.P + pa;
        const float* fb = fimg + (kc0 + kb) * a.P + pb;
#pragma unroll 2
        for (int kk = 0; kk < KW; kk += 4) {
            const float a0 = ldg_stream_f(fa + (kk + 0) * a.P), a1 = ldg_stream_f(fa + (kk + 1) * a.P);
            const float a2 = ldg_stream_f(fa + (kk + 2) * a.P), a3 = ldg_stream_f(fa + (kk + 3) * a.P);
            const float b0 = ldg_stream_f(fb + (kk + 0) * a.P), b1 = ldg_stream_f(fb + (kk + 1) * a.P);
            const float b2 = ldg_stream_f(fb + (kk + 2) * a.P), b3 = ldg_stream_f(fb + (kk + 3) * a.P);
            const f32x2 A01 = pack2(a0, a1), A23 = pack2(a2, a3), B01 = pack2(b0, b1), B23 = pack2(b2, b3);
#pragma unroll
            for (int c = 0; c < CT; ++c) {
                const ulonglong2 p = *reinterpret_cast<const ulonglong2*>(ps + c * PS + kb + kk);
                f32x2 d;
                d = sub2(p.x, A01); accA[c] = fma2(d, d, accA[c]);
                d = sub2(p.y, A23); accA[c] = fma2(d, d, accA[c]);
                d = sub2(p.x, B01); accB[c] = fma2(d, d, accB[c]);
                d = sub2(p.y, B23); accB[c] = fma2(d, d, accB[c]);
            }
        }
    }
    __syncthreads();
    float* red = ps;  // [8][CT][64]
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
    if (tid >= kDTile) return;
    const int64_t p = p0 + tid;
    if (p >= a.P) return;
    float d[CT];
    float dmin = INFINITY;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) {
            d[c] = dsm[c * kDTile + tid];
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
    float best = 0.f, rsum = 0.f;
    int lab = 0;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) {
            const float r = d[c] * a.soft[obase + c * a.P];
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

template <int CT>
static int launch_distance(const DistArgs& a, cudaStream_t st) {
    const int PS = a.KC + 4;
    const int main_floats = (CT * PS > kDWarps * CT * kDTile) ? CT * PS : kDWarps * CT * kDTile;
    const size_t smem = sizeof(float) * (static_cast<size_t>(main_floats) + CT * kDTile);
    cudaError_t e = cudaFuncSetAttribute(distance_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return cuda_status(e, "tss_proto_distance: smem attribute");
    const int64_t tiles = a.N * ((a.P + kDTile - 1) / kDTile);
    TSS_REQUIRE(tiles < (1ll << 31), TSS_E_ARG, "tss_proto_distance: too many tiles");
    distance_kernel<CT><<<static_cast<unsigned>(tiles), kDThreads, smem, st>>>(a);
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
    // K chunk staged per pass: the largest 32*g <= 512 dividing K
    const int64_t m = channels / 32;
    int g = 1;
    for (int cand = 16; cand >= 1; --cand)
        if (m % cand == 0) { g = cand; break; }
    a.KC = 32 * g;
    cudaStream_t st = as_stream(stream);
    if (num_classes <= 4) return launch_distance<4>(a, st);
    if (num_classes <= 8) return launch_distance<8>(a, st);
    if (num_classes <= 13) return launch_distance<13>(a, st);
    if (num_classes <= 16) return launch_distance<16>(a, st);
    if (num_classes <= 19) return launch_distance<19>(a, st);
    return launch_distance<20>(a, st);
}
