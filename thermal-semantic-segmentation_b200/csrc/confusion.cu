// A14 confusion matrix (utils/eval_tools.py:4-9, :49-51).  HBM-bound integer work: 2 B/px for the
// uint8 path, 16 B/px for the reference's int64 arrays.
//
// Two kernels:
//  * confusion_atomic<T>: any label dtype.  Persistent grid (multiple of 148 CTAs); every warp owns a
//    private C*C uint32 histogram in shared memory and bumps it with shared-memory atomics; one global
//    64-bit atomic per non-zero bin per CTA at the end.
//  * confusion_packed<T>: 16 pixels per lane and round, 128-bit loads (1 per array for uint8, 8 for int64).  Shared-memory atomics cannot keep
//    up with 2 B/px at HBM speed (~12 px/clk/SM), so each THREAD owns a private histogram of packed
//    8-bit counters laid out [word][lane] (bank == lane: conflict-free, no atomics), flushed into the
//    CTA histogram every <= 240 pixels per thread, before any counter can wrap.
#include "common.cuh"

namespace tss {

constexpr int kHistThreads = 256;

template <typename T>
__global__ void __launch_bounds__(kHistThreads)
confusion_atomic(const T* __restrict__ pred, const T* __restrict__ gt, int64_t n, int C,
                 unsigned long long* __restrict__ hist, unsigned long long* __restrict__ bad) {
    extern __shared__ unsigned int sh[];  // [warps][C*C]
    const int bins = C * C;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < bins * (kHistThreads / 32); i += kHistThreads) sh[i] = 0;
    __syncthreads();
    unsigned int* mine = sh + warp * bins;
    unsigned int nbad = 0;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kHistThreads;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * kHistThreads + threadIdx.x; i < n; i += stride) {
        long long g = static_cast<long long>(gt[i]);
        long long p = static_cast<long long>(pred[i]);
        if (g >= 0 && g < C) {
            if (p >= 0 && p < C) atomicAdd(&mine[static_cast<int>(g) * C + static_cast<int>(p)], 1u);
            else ++nbad;
        }
    }
    __syncthreads();
    for (int b = threadIdx.x; b < bins; b += kHistThreads) {
        unsigned long long s = 0;
#pragma unroll
        for (int w = 0; w < kHistThreads / 32; ++w) s += sh[w * bins + b];
        if (s) atomicAdd(&hist[b], s);
    }
    nbad = __reduce_add_sync(0xffffffffu, nbad);
    if (bad != nullptr && (threadIdx.x & 31) == 0 && nbad) atomicAdd(bad, static_cast<unsigned long long>(nbad));
}

// ---------------------------------------------------------------------------------------------
// Packed private counters.  CP = C + 1: row CP-1 collects ignored ground truth (gt >= C), column
// CP-1 collects out-of-range predictions.  bin = gt' * CP + pred', four 8-bit counters per word.
constexpr int kPackedWarps = 8;
constexpr int kPackedIters = 15;  // 15 * 16 px = 240 <= 255 increments per flush

__device__ __forceinline__ unsigned int min_u16x2(unsigned int a, unsigned int b) {
    unsigned int r;
    asm("min.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
}

__device__ __forceinline__ void bump(unsigned int* lane_base, unsigned int key) {
    // word (key >> 2) of this lane's column, byte (key & 3)
    unsigned int* w = lane_base + ((key >> 2) << 5);
    *w += 1u << ((key << 3) & 31u);
}

__device__ __forceinline__ void count4(unsigned int* lane_base, unsigned int g4, unsigned int p4, unsigned int cp,
                                       unsigned int cmax2) {
    // bytes -> two u16x2 pairs, clamp to CP-1, key = g * CP + p on both halves at once
    unsigned int g01 = min_u16x2(__byte_perm(g4, 0, 0x4140), cmax2);
    unsigned int g23 = min_u16x2(__byte_perm(g4, 0, 0x4342), cmax2);
    unsigned int p01 = min_u16x2(__byte_perm(p4, 0, 0x4140), cmax2);
    unsigned int p23 = min_u16x2(__byte_perm(p4, 0, 0x4342), cmax2);
    unsigned int k01 = g01 * cp + p01;  // fields stay below 2^16: (CP-1)*CP + CP-1 < 1089
    unsigned int k23 = g23 * cp + p23;
    bump(lane_base, k01 & 0xffffu);
    bump(lane_base, k01 >> 16);
    bump(lane_base, k23 & 0xffffu);
    bump(lane_base, k23 >> 16);
}

// 16 pixels of one lane as four packed-byte quads, values clamped to <= C (anything outside [0, C) becomes C).
// Every load instruction of the warp covers 512 contiguous bytes (lane l takes bytes [16 l, 16 l + 16)), so wider
// label types take several loads per 16 pixels: `p` points at the first element of the warp's 512-pixel block.
template <typename T>
__device__ __forceinline__ void load_quads(const T* p, int lane, unsigned int C, unsigned int (&q)[4]);
template <>
__device__ __forceinline__ void load_quads<uint8_t>(const uint8_t* p, int lane, unsigned int, unsigned int (&q)[4]) {
    const uint4 x = ldg_stream_u4(p + lane * 16);   // clamped later by count4's min
    q[0] = x.x; q[1] = x.y; q[2] = x.z; q[3] = x.w;
}
__device__ __forceinline__ unsigned int clamp_pack4(unsigned int a, unsigned int b, unsigned int c, unsigned int d, unsigned int C) {
    return min(a, C) | (min(b, C) << 8) | (min(c, C) << 16) | (min(d, C) << 24);   // unsigned min: negatives -> C
}
template <>
__device__ __forceinline__ void load_quads<int32_t>(const int32_t* p, int lane, unsigned int C, unsigned int (&q)[4]) {
    uint4 x[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) x[k] = ldg_stream_u4(p + k * 128 + lane * 4);
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = clamp_pack4(x[k].x, x[k].y, x[k].z, x[k].w, C);
}
template <>
__device__ __forceinline__ void load_quads<int64_t>(const int64_t* p, int lane, unsigned int C, unsigned int (&q)[4]) {
    uint4 x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = ldg_stream_u4(p + k * 64 + lane * 2);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint4 a = x[2 * k], b = x[2 * k + 1];   // two int64 each: {lo, hi, lo, hi}
        q[k] = clamp_pack4(a.y ? C : a.x, a.w ? C : a.z, b.y ? C : b.x, b.w ? C : b.z, C);
    }
}

template <typename T>
__global__ void __launch_bounds__(kPackedWarps * 32)
confusion_packed(const T* __restrict__ pred, const T* __restrict__ gt, int64_t n_vec, int C, int iters,
                 unsigned long long* __restrict__ hist, unsigned long long* __restrict__ bad) {
    // n_vec = number of 16-pixel vectors, a multiple of 32 (both pointers 16-byte aligned; the tail goes to the
    // atomic kernel)
    extern __shared__ unsigned int sh[];
    const int CP = C + 1;
    const int bins = CP * CP;
    const int words = (bins + 3) >> 2;
    unsigned int* cta_hist = sh;                                   // [bins] uint32
    unsigned int* priv = sh + ((bins + 31) & ~31);                 // [warps][words][32]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned int* warp_base = priv + warp * words * 32;
    unsigned int* lane_base = warp_base + lane;
    for (int i = threadIdx.x; i < bins; i += blockDim.x) cta_hist[i] = 0;
    for (int i = lane; i < words * 32; i += 32) warp_base[i] = 0;
    __syncthreads();
    const unsigned int cmax2 = static_cast<unsigned int>(C) * 0x00010001u;
    // chunk = `iters` consecutive warp-wide rounds of 512 pixels; chunks are dealt round-robin to all warps
    const int64_t warps_total = static_cast<int64_t>(gridDim.x) * kPackedWarps;
    const int64_t chunk_vecs = static_cast<int64_t>(iters) * 32;  // iters <= kPackedIters: no counter can wrap
    for (int64_t chunk = static_cast<int64_t>(blockIdx.x) * kPackedWarps + warp;; chunk += warps_total) {
        const int64_t v0 = chunk * chunk_vecs;
        if (v0 >= n_vec) break;
#pragma unroll 1
        for (int it = 0; it < iters; ++it) {
            const int64_t vb = v0 + static_cast<int64_t>(it) * 32;   // first vector of this round (warp-uniform)
            if (vb < n_vec) {
                unsigned int pq[4], gq[4];
                load_quads<T>(pred + vb * 16, lane, static_cast<unsigned int>(C), pq);
                load_quads<T>(gt + vb * 16, lane, static_cast<unsigned int>(C), gq);
#pragma unroll
                for (int k = 0; k < 4; ++k) count4(lane_base, gq[k], pq[k], CP, cmax2);
            }
        }
        // flush this warp's 32 private columns into the CTA histogram; lane l sums word (l + 32 r) over the
        // 32 columns in rotated order so that every access hits 32 distinct banks
        __syncwarp();
        for (int w0 = 0; w0 < words; w0 += 32) {
            const int w = w0 + lane;
            if (w < words) {
                unsigned int lo = 0, hi = 0;  // two 16-bit fields each: bytes {0,2} and {1,3}
                // 128-bit reads of four columns at a time; a quarter warp (8 lanes, the unit a 128-bit shared access
                // is served in) reads eight different 16-byte chunks = all 32 banks
                uint4* rowq = reinterpret_cast<uint4*>(warp_base + w * 32);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint4* q = rowq + ((lane + j) & 7);
                    const uint4 x = *q;
                    *q = make_uint4(0u, 0u, 0u, 0u);
                    lo += (x.x & 0x00ff00ffu) + (x.y & 0x00ff00ffu);
                    hi += __byte_perm(x.x, 0, 0x4341) + __byte_perm(x.y, 0, 0x4341);   // bytes {1, 3} into the 16-bit fields
                    lo += (x.z & 0x00ff00ffu) + (x.w & 0x00ff00ffu);
                    hi += __byte_perm(x.z, 0, 0x4341) + __byte_perm(x.w, 0, 0x4341);
                }
                const int b = w << 2;
                if (lo & 0xffffu) atomicAdd(&cta_hist[b], lo & 0xffffu);
                if (b + 1 < bins && (hi & 0xffffu)) atomicAdd(&cta_hist[b + 1], hi & 0xffffu);
                if (b + 2 < bins && (lo >> 16)) atomicAdd(&cta_hist[b + 2], lo >> 16);
                if (b + 3 < bins && (hi >> 16)) atomicAdd(&cta_hist[b + 3], hi >> 16);
            }
        }
        __syncwarp();
    }
    __syncthreads();
    unsigned long long nbad = 0;
    for (int b = threadIdx.x; b < bins; b += blockDim.x) {
        const unsigned int s = cta_hist[b];
        if (!s) continue;
        const int g = b / CP, p = b - g * CP;
        if (g == C) continue;  // ignored ground truth
        if (p == C) nbad += s;
        else atomicAdd(&hist[g * C + p], static_cast<unsigned long long>(s));
    }
    if (bad != nullptr && nbad) atomicAdd(bad, nbad);
}

}  // namespace tss

extern "C" int tss_confusion(const void* pred, const void* gt, int dtype, int64_t n, int num_classes,
                             int64_t* hist, uint64_t* bad_keys, int variant, void* stream) {
    using namespace tss;
    TSS_REQUIRE(hist != nullptr && n >= 0, TSS_E_ARG, "tss_confusion: null hist or negative n");
    TSS_REQUIRE(n == 0 || (pred != nullptr && gt != nullptr), TSS_E_ARG, "tss_confusion: null input");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES,
                "tss_confusion: num_classes %d outside [1, %d]", num_classes, TSS_MAX_CLASSES);
    TSS_REQUIRE(dtype == TSS_U8 || dtype == TSS_I32 || dtype == TSS_I64, TSS_E_ARG, "tss_confusion: bad dtype %d", dtype);
    TSS_REQUIRE(variant >= TSS_HIST_AUTO && variant <= TSS_HIST_PACKED, TSS_E_ARG, "tss_confusion: bad variant %d", variant);
    if (n == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    auto* h = reinterpret_cast<unsigned long long*>(hist);
    auto* b = reinterpret_cast<unsigned long long*>(bad_keys);
    const int C = num_classes;
    const size_t sh_atomic = static_cast<size_t>(C) * C * (kHistThreads / 32) * sizeof(unsigned int);
    auto grid_for = [](int64_t items, int per_cta, int ctas_per_sm) {
        int64_t need = (items + per_cta - 1) / per_cta;
        int64_t cap = static_cast<int64_t>(kNumSMs) * ctas_per_sm;
        return static_cast<int>(need < cap ? (need < 1 ? 1 : need) : cap);
    };
    int64_t done = 0;
    if (variant != TSS_HIST_ATOMIC && n >= 16 * 32 * kPackedIters &&
        (reinterpret_cast<uintptr_t>(pred) & 15) == 0 && (reinterpret_cast<uintptr_t>(gt) & 15) == 0) {
        const int CP = C + 1, bins = CP * CP, words = (bins + 3) / 4;
        const size_t sh = (static_cast<size_t>((bins + 31) & ~31) + static_cast<size_t>(kPackedWarps) * words * 32) * 4;
        cudaError_t e = cudaSuccess;
        if (dtype == TSS_U8) e = cudaFuncSetAttribute(confusion_packed<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        else if (dtype == TSS_I32) e = cudaFuncSetAttribute(confusion_packed<int32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        else e = cudaFuncSetAttribute(confusion_packed<int64_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return cuda_status(e, "tss_confusion: smem attribute");
        const int64_t n_vec = (n / 512) * 32;   // whole 512-pixel warp rounds
        int ctas_per_sm = static_cast<int>((220 * 1024) / (sh + 1024));
        if (ctas_per_sm > 8) ctas_per_sm = 8;
        if (ctas_per_sm < 1) ctas_per_sm = 1;
        // chunk length: as long as possible (fewer flushes) while one wave of resident warps covers the input
        const int64_t slots = static_cast<int64_t>(kNumSMs) * ctas_per_sm * kPackedWarps;
        int64_t iters = (n_vec + 32 * slots - 1) / (32 * slots);
        if (iters > kPackedIters) iters = kPackedIters;
        if (iters < 1) iters = 1;
        const int grid = grid_for(n_vec, kPackedWarps * 32 * static_cast<int>(iters), ctas_per_sm);
        if (dtype == TSS_U8)
            confusion_packed<uint8_t><<<grid, kPackedWarps * 32, sh, st>>>(static_cast<const uint8_t*>(pred), static_cast<const uint8_t*>(gt), n_vec, C, static_cast<int>(iters), h, b);
        else if (dtype == TSS_I32)
            confusion_packed<int32_t><<<grid, kPackedWarps * 32, sh, st>>>(static_cast<const int32_t*>(pred), static_cast<const int32_t*>(gt), n_vec, C, static_cast<int>(iters), h, b);
        else
            confusion_packed<int64_t><<<grid, kPackedWarps * 32, sh, st>>>(static_cast<const int64_t*>(pred), static_cast<const int64_t*>(gt), n_vec, C, static_cast<int>(iters), h, b);
        TSS_LAUNCH_CHECK("tss_confusion(packed)");
        done = n_vec * 16;
    }
    if (done < n) {
        const int64_t rest = n - done;
        const int grid = grid_for(rest, kHistThreads * 8, 8);
        switch (dtype) {
            case TSS_U8:
                confusion_atomic<uint8_t><<<grid, kHistThreads, sh_atomic, st>>>(
                    static_cast<const uint8_t*>(pred) + done, static_cast<const uint8_t*>(gt) + done, rest, C, h, b);
                break;
            case TSS_I32:
                confusion_atomic<int32_t><<<grid, kHistThreads, sh_atomic, st>>>(
                    static_cast<const int32_t*>(pred) + done, static_cast<const int32_t*>(gt) + done, rest, C, h, b);
                break;
            default:
                confusion_atomic<int64_t><<<grid, kHistThreads, sh_atomic, st>>>(
                    static_cast<const int64_t*>(pred) + done, static_cast<const int64_t*>(gt) + done, rest, C, h, b);
                break;
        }
        TSS_LAUNCH_CHECK("tss_confusion(atomic)");
    }
    return TSS_OK;
}
