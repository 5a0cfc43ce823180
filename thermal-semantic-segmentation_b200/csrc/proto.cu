// A5/A6: per-(image, class) segmented sums of the 1/8-stride features by argmax label, and the
// prototype update (cal_prototype.py:118-169, self_training.py:176-227).
//
// Data layout in HBM: feat [N, K, P] fp32 exactly as the network leaves it (P = h*w is ODD for every
// DeepLab shape, so channel rows are only 4-byte aligned).  The reference re-reads feat up to C times
// and materialises a [K,h,w] product per (image, class); here feat is streamed ONCE:
//
//   proto_label_kernel   logits -> argmax label per pixel, class counts, and per 512-pixel tile a
//                        class-sorted list of pixel offsets in GROUPS OF EIGHT: every class segment is padded
//                        to a multiple of eight entries with the offset of a slot that always holds 0.0f, so a
//                        group belongs to exactly one class and the accumulate kernel has no edge handling.
//   proto_accum_kernel   persistent, one CTA per SM.  A producer warp stages tiles of 32 channel rows x
//                        512 pixels into shared memory with 1-D bulk async copies (TMA engine, UBLKCP)
//                        completing on an mbarrier ring; every lane copies the 16-byte-aligned superset
//                        of its (misaligned) row, rows are skewed so that "lane = channel" reads are
//                        bank-conflict free.  NC consumer warps (16 when the kernel has the SM to itself, 4 when
//                        it shares the SM with the pseudo-label kernel) each take a contiguous run of the tile's
//                        groups: one 128-bit shared load brings eight pixel offsets (warp-uniform), eight
//                        unconditional shared loads + an add tree, and the running sum goes to ONE register
//                        accumulator per class when the class changes -- no atomics, no select chains.
//                        Global fp32 reductions happen once per class per (image, channel group) per warp.
#include "common.cuh"
#include <mutex>

namespace tss {

constexpr int kPT = TSS_PROTO_TILE;        // pixels per tile
constexpr int kKT = 32;                    // channels per tile: one per lane
constexpr int kRowStride = kPT + 40;       // floats; head misalignment (3) + 16 B round-up (3) + zero slots (4) + skew (28)
constexpr int kPadOff = (kPT + 7) * 4;     // byte offset (from a lane's first pixel) of its always-zero slot: past every
                                           // byte a bulk copy can write (<= 3 + 512 + 3 floats from the aligned start)
constexpr int kIdxMax = kPT + 8 * TSS_MAX_CLASSES;   // sum_c roundup8(count_c) <= 512 + 7 C
constexpr int kGroupsMax = kIdxMax / 8;
constexpr int kStages = 3;

static_assert(kRowStride % 4 == 0, "rows must start 16-byte aligned");

struct TileRecord {
    uint16_t idx[kIdxMax];      // BYTE offsets (4 x pixel index, or kPadOff) of the tile's pixels, sorted by class
    uint8_t gcls[kGroupsMax];   // class of every group of eight entries
    uint16_t cnt[32];           // gated pixel count per class in the tile (the divisor)
    uint16_t call[32];          // ungated one-hot count per class in the tile (the "< 10 pixels" rule)
    uint16_t groups;            // groups in use
    uint16_t pad_[7];
};
constexpr int kRecBytes = sizeof(TileRecord);
static_assert(kRecBytes % 16 == 0, "records are bulk-copied");
static_assert(kRecBytes == TSS_PROTO_RECORD_BYTES, "header constant");

// ------------------------------------------------------------------------------------------------
template <int CT>
__global__ void __launch_bounds__(256, 5)
proto_label_kernel(const float* __restrict__ logits, const float* __restrict__ labels_val, int C, int64_t P,
                   int overflow, int use_gate, float thresh, uint8_t* __restrict__ seg_label,
                   TileRecord* __restrict__ records, float* __restrict__ zero_buf, int64_t zero_elems) {
    __shared__ int wcnt[16][TSS_MAX_CLASSES + 1];
    __shared__ int call[TSS_MAX_CLASSES], cgate[TSS_MAX_CLASSES];
    __shared__ int boff[TSS_MAX_CLASSES + 2];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t tile = blockIdx.x, n = blockIdx.y;
    const int64_t tiles = gridDim.x;
    const int64_t p0 = tile * kPT;
    // zero the segmented-sum buffer the accumulate kernel adds into (saves a memset node on the stream)
    for (int64_t i = (n * tiles + tile) * 256 + tid; i < zero_elems; i += tiles * gridDim.y * 256) zero_buf[i] = 0.f;
    for (int i = tid; i < 16 * (TSS_MAX_CLASSES + 1); i += 256) (&wcnt[0][0])[i] = 0;
    if (tid < TSS_MAX_CLASSES) { call[tid] = 0; cgate[tid] = 0; }
    __syncthreads();
    int bucket[2], rank[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int pl = r * 256 + tid;
        const int64_t p = p0 + pl;
        int lab = 255, hot_class = -1;
        if (p < P) {
            const float* x = logits + n * C * P + p;
            float v[CT];
#pragma unroll
            for (int c = 0; c < CT; ++c) v[c] = ldg_stream_f(x + static_cast<int64_t>(min(c, C - 1)) * P);  // all in flight
            float m = v[0];
            int t = 0;
#pragma unroll
            for (int c = 1; c < CT; ++c) {
                const bool gt = c < C && v[c] > m;   // strict: ties keep the lowest class (torch argmax)
                m = gt ? v[c] : m;
                t = gt ? c : t;
            }
            bool hot = true;
            if (labels_val != nullptr) {
                const float lv = labels_val[n * P + p];
                const int lvc = (lv < static_cast<float>(C)) ? static_cast<int>(lv) : overflow;
                hot = (lvc == t);
            }
            bool gate = true;
            if (use_gate) {
                float s = 0.f;
                const float mb = m * kLog2e;
#pragma unroll
                for (int c = 0; c < CT; ++c)
                    if (c < C) s += ex2_approx(fmaf(v[c], kLog2e, -mb));
                gate = __fdiv_rn(1.0f, s) >= thresh;
            }
            if (hot) {
                hot_class = t;
                if (gate) lab = t;
            }
            seg_label[n * P + p] = static_cast<uint8_t>(lab);
        }
        bucket[r] = (lab == 255) ? C : lab;
        const unsigned int peers = __match_any_sync(0xffffffffu, bucket[r]);
        rank[r] = __popc(peers & ((1u << lane) - 1u));
        if (rank[r] == 0) wcnt[r * 8 + warp][bucket[r]] = __popc(peers);
        if (use_gate || labels_val != nullptr) {
            // ungated one-hot count (the "< 10" rule) differs from the gated one: count it separately,
            // one shared atomic per (warp, class) instead of one per pixel
            const unsigned int hp = __match_any_sync(0xffffffffu, hot_class);
            if (hot_class >= 0 && __popc(hp & ((1u << lane) - 1u)) == 0) atomicAdd(&call[hot_class], __popc(hp));
        }
    }
    __syncthreads();
    // totals per bucket, then class segments padded to a multiple of eight entries (bucket C is not stored)
    if (tid <= C) {
        int tot = 0;
        for (int s = 0; s < 16; ++s) tot += wcnt[s][tid];
        boff[tid + 1] = tot;
    }
    __syncthreads();
    __shared__ int pstart[TSS_MAX_CLASSES + 1];
    if (tid == 0) {
        int run = 0;
        for (int b = 0; b < C; ++b) {
            pstart[b] = run;
            run += (boff[b + 1] + 7) & ~7;
        }
        pstart[C] = run;
    }
    __syncthreads();
    TileRecord* rec = records + n * tiles + tile;
    if (tid < C) {
        int run = pstart[tid];
        for (int s = 0; s < 16; ++s) {
            const int c = wcnt[s][tid];
            wcnt[s][tid] = run;
            run += c;
        }
        for (; run < pstart[tid + 1]; ++run) rec->idx[run] = static_cast<uint16_t>(kPadOff);
        for (int g = pstart[tid] >> 3; g < (pstart[tid + 1] >> 3); ++g) rec->gcls[g] = static_cast<uint8_t>(tid);
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 2; ++r)
        if (bucket[r] < C) rec->idx[wcnt[r * 8 + warp][bucket[r]] + rank[r]] = static_cast<uint16_t>((r * 256 + tid) * 4);
    if (tid < 32) {
        const int g = tid < C ? boff[tid + 1] : 0;  // gated count of this class in the tile
        rec->cnt[tid] = static_cast<uint16_t>(g);
        rec->call[tid] = static_cast<uint16_t>(tid < C ? ((use_gate || labels_val != nullptr) ? call[tid] : g) : 0);
    }
    if (tid == 0) rec->groups = static_cast<uint16_t>(pstart[C] >> 3);
}

// ------------------------------------------------------------------------------------------------
struct AccSmem {
    // dynamic shared memory carve-up (byte offsets)
    static constexpr int rows_bytes = kStages * kKT * kRowStride * 4;
    static constexpr int rec_off = rows_bytes;
    static constexpr int bar_off = rec_off + kStages * kRecBytes;
    static constexpr int total = bar_off + 2 * kStages * 8 + 16;
};
static_assert(AccSmem::total <= 227 * 1024, "accumulate kernel shared memory");

__device__ __forceinline__ float lds_f32(uint32_t addr) {  // unconditional shared load (never sunk into a branch)
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ int row_skew_words(int lane) {
    // B_lane = lane*kRowStride + 4*g: choose g so that (B_lane / 4) mod 8 == (lane >> 2) & 7.  With the head
    // misalignment o_lane = (e0 + lane*P) & 3 (distinct for lane & 3 when P is odd) the 32 lanes of a
    // "same pixel, lane = channel" read hit 32 distinct banks.
    constexpr int s = kRowStride / 4;
    return 4 * (((lane >> 2) - lane * s) & 7);
}

// racc[c] += t for a warp-uniform run-time class c: a switch (one indirect branch, no divergence) instead of a
// chain of CT predicated adds
template <int CT>
__device__ __forceinline__ void add_to_class(float (&racc)[CT], int c, float t) {
#define TSS_ACC_CASE(i) \
    case i:             \
        if (i < CT) racc[i < CT ? i : 0] += t; \
        break;
    switch (c) {
        TSS_ACC_CASE(0)
        TSS_ACC_CASE(1)
        TSS_ACC_CASE(2)
        TSS_ACC_CASE(3)
        TSS_ACC_CASE(4)
        TSS_ACC_CASE(5)
        TSS_ACC_CASE(6)
        TSS_ACC_CASE(7)
        TSS_ACC_CASE(8)
        TSS_ACC_CASE(9)
        TSS_ACC_CASE(10)
        TSS_ACC_CASE(11)
        TSS_ACC_CASE(12)
        TSS_ACC_CASE(13)
        TSS_ACC_CASE(14)
        TSS_ACC_CASE(15)
        TSS_ACC_CASE(16)
        TSS_ACC_CASE(17)
        TSS_ACC_CASE(18)
        TSS_ACC_CASE(19)
        TSS_ACC_CASE(20)
        TSS_ACC_CASE(21)
        TSS_ACC_CASE(22)
        TSS_ACC_CASE(23)
        TSS_ACC_CASE(24)
        TSS_ACC_CASE(25)
        TSS_ACC_CASE(26)
        TSS_ACC_CASE(27)
        TSS_ACC_CASE(28)
        TSS_ACC_CASE(29)
        TSS_ACC_CASE(30)
        TSS_ACC_CASE(31)
        default: break;
    }
#undef TSS_ACC_CASE
}

// NC = consumer warps.  The register cap is 96 per thread in every form (minBlocks is only the cap: shared memory
// holds one CTA per SM), so that NC = 4 (160 threads, 15 K registers) fits beside three 4-warp CTAs of the
// pseudo-label kernel on the same SM.
template <int NC>
struct AccCfg {
    static constexpr int kThreads = (NC + 1) * 32;
    static constexpr int kMinBlocks = 65536 / (kThreads * 96) < 1 ? 1 : 65536 / (kThreads * 96);
};

template <int CT, int NC>
__global__ void __launch_bounds__(AccCfg<NC>::kThreads, AccCfg<NC>::kMinBlocks)
proto_accum_kernel(const float* __restrict__ feat, const TileRecord* __restrict__ records, int64_t N, int64_t K,
                   int64_t P, int C, float* __restrict__ sums) {
    extern __shared__ __align__(128) unsigned char smem[];
    float* rows = reinterpret_cast<float*>(smem);
    unsigned char* recs = smem + AccSmem::rec_off;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + AccSmem::bar_off);
    uint64_t* empty = full + kStages;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // the always-zero slots (kPadOff) live in the staged rows: clear the whole row area once; bulk copies never
    // reach those slots afterwards
    for (int i = tid; i < AccSmem::rows_bytes / 16; i += AccCfg<NC>::kThreads)
        reinterpret_cast<float4*>(rows)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int64_t T = (P + kPT - 1) / kPT;
    const int64_t KG = K / kKT;
    const int64_t jobs = N * KG * T;
    const int64_t j0 = jobs * blockIdx.x / gridDim.x, j1 = jobs * (blockIdx.x + 1) / gridDim.x;
    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 32);
            mbar_init(&empty[s], NC);
        }
        mbar_fence_init();
    }
    // generic-proxy stores (the zero fill) before async-proxy writes (bulk copies) to the same shared memory
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int64_t total_elems = N * K * P;

    if (warp == NC) {
        // ===== producer warp: lane l stages channel row l of every tile job
        int stage = 0;
        uint32_t phase = 0;
        // (tile, channel group, image) walked incrementally: a 64-bit division per tile and warp is ~100 instructions
        int64_t t = j0 % T, u0 = j0 / T;
        int64_t kg = u0 % KG, n = u0 / KG;
        for (int64_t j = j0; j < j1; ++j, ++t) {
            if (t == T) {
                t = 0;
                if (++kg == KG) { kg = 0; ++n; }
            }
            const int64_t p0 = t * kPT;
            const int len = static_cast<int>(min(static_cast<int64_t>(kPT), P - p0));
            mbar_wait(&empty[stage], phase ^ 1u);
            const int64_t e0 = (n * K + kg * kKT + lane) * P + p0;
            const int o = static_cast<int>(e0 & 3);
            const int64_t ea = e0 - o;
            uint32_t nbytes = static_cast<uint32_t>(((o + len) * 4 + 15) & ~15);
            int patch = 0;
            if ((ea + nbytes / 4) > total_elems) {  // never read past the end of the tensor
                nbytes = static_cast<uint32_t>((total_elems - ea) * 4) & ~15u;
                patch = (o + len) - static_cast<int>(nbytes / 4);
            }
            float* dst = rows + (stage * kKT + lane) * kRowStride + row_skew_words(lane);
            for (int i = 0; i < patch; ++i) dst[nbytes / 4 + i] = feat[ea + nbytes / 4 + i];
            const uint32_t tx = nbytes + (lane == 0 ? kRecBytes : 0);
            mbar_arrive_expect_tx(&full[stage], tx);
            if (nbytes) bulk_g2s(dst, feat + ea, nbytes, &full[stage]);
            if (lane == 0) bulk_g2s(recs + stage * kRecBytes, records + n * T + t, kRecBytes, &full[stage]);
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
    } else {
        // ===== consumer warps: class-sorted walk, lane = channel.  Warp w takes the contiguous run
        // [G w / NC, G (w+1) / NC) of the tile's G groups; every class has its own REGISTER accumulator.  When the
        // (image, channel group) changes, every warp adds its registers straight to the global sums with
        // fire-and-forget reductions (RED.ADD.F32 resolves in L2, the SM does not wait).
        float racc[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) racc[c] = 0.f;
        int stage = 0;
        uint32_t phase = 0;
        int64_t cur_unit = -1;
        int64_t cur_n = 0, cur_kg = 0;
        auto flush = [&](int64_t n, int64_t kg) {
            float* out = sums + (n * C) * K + kg * kKT + lane;
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (c < C) {
                    if (racc[c] != 0.f) atomicAdd(out + c * K, racc[c]);
                    racc[c] = 0.f;
                }
        };
        int64_t t = j0 % T, u = j0 / T;
        int64_t kg = u % KG, n = u / KG;
        for (int64_t j = j0; j < j1; ++j, ++t) {
            if (t == T) {
                t = 0;
                ++u;
                if (++kg == KG) { kg = 0; ++n; }
            }
            if (u != cur_unit) {
                if (cur_unit >= 0) flush(cur_n, cur_kg);
                cur_unit = u;
                cur_n = n;
                cur_kg = kg;
            }
            const int64_t e0 = (n * K + kg * kKT + lane) * P + t * kPT;
            const float* row = rows + (stage * kKT + lane) * kRowStride + row_skew_words(lane) + static_cast<int>(e0 & 3);
            const TileRecord* rec = reinterpret_cast<const TileRecord*>(recs + stage * kRecBytes);
            mbar_wait(&full[stage], phase);
#ifndef TSS_PROTO_NOCONSUME
            const int G = rec->groups;
            const int g0 = (G * warp) / NC, g1 = (G * (warp + 1)) / NC;
            const uint32_t row_addr = smem_u32(row);
            int cur = -1;
            float run = 0.f;
            for (int g = g0; g < g1; ++g) {  // warp-uniform
                const int c = rec->gcls[g];
                if (c != cur) {
                    if (cur >= 0) add_to_class<CT>(racc, cur, run);
                    run = 0.f;
                    cur = c;
                }
                // ONE 128-bit shared load brings the eight byte offsets; every offset of a record is a valid offset
                // into the staged row (a pixel of the tile, or the zero slot)
                const uint4 iq = *reinterpret_cast<const uint4*>(&rec->idx[8 * g]);
                float v[8];
                v[0] = lds_f32(row_addr + (iq.x & 0xffffu));
                v[1] = lds_f32(row_addr + (iq.x >> 16));
                v[2] = lds_f32(row_addr + (iq.y & 0xffffu));
                v[3] = lds_f32(row_addr + (iq.y >> 16));
                v[4] = lds_f32(row_addr + (iq.z & 0xffffu));
                v[5] = lds_f32(row_addr + (iq.z >> 16));
                v[6] = lds_f32(row_addr + (iq.w & 0xffffu));
                v[7] = lds_f32(row_addr + (iq.w >> 16));
                run += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
            }
            if (cur >= 0) add_to_class<CT>(racc, cur, run);
#endif
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (cur_unit >= 0) flush(cur_n, cur_kg);
    }
}

// ------------------------------------------------------------------------------------------------
// per-(image, class) pixel counts = sums over the image's tile records (no atomics, no zero-initialised buffers)
__device__ __forceinline__ void tile_counts(const TileRecord* __restrict__ rec, int64_t tiles, int c, int lane,
                                            int& gated, int& all) {
    int g = 0, a = 0;
    for (int64_t t = lane; t < tiles; t += 32) {
        g += rec[t].cnt[c];
        a += rec[t].call[c];
    }
    gated = __reduce_add_sync(0xffffffffu, g);
    all = __reduce_add_sync(0xffffffffu, a);
}

// Destinations in other GPUs' memory (peer-mapped, e.g. symmetric-memory buffers): the multi-GPU prototype
// exchange is the finalize kernel's own store stream, not a separate collective.
constexpr int kMaxPeers = 7;
struct PeerList {
    float* dst[kMaxPeers];   // [N,C,K] mean vectors
    float* sum[kMaxPeers];   // [N,C] their sums over K (NULL: not sent)
    int n;
};

__global__ void __launch_bounds__(256)
proto_finalize_kernel(const float* __restrict__ sums, const TileRecord* __restrict__ records, int64_t tiles, int C,
                      int64_t K, int min_pixels, int32_t* __restrict__ cnt_all, int32_t* __restrict__ cnt_gated,
                      float* __restrict__ vectors, uint8_t* __restrict__ valid, float* __restrict__ vector_sum,
                      const PeerList peers) {
    __shared__ float part[8];
    __shared__ int cnt[2];
    const int64_t nc = blockIdx.x;  // n*C + c
    const int64_t n = nc / C;
    const int c = static_cast<int>(nc - n * C);
    if (threadIdx.x < 32) {
        int g, a;
        tile_counts(records + n * tiles, tiles, c, threadIdx.x, g, a);
        if (threadIdx.x == 0) { cnt[0] = g; cnt[1] = a; }
    }
    __syncthreads();
    const int cg = cnt[0], ca = cnt[1];
    const bool ok = cg > 0 && ca >= min_pixels;
    const float denom = static_cast<float>(cg);
    float s = 0.f;
    for (int64_t k = threadIdx.x; k < K; k += 256) {
        const float v = ok ? __fdiv_rn(sums[nc * K + k], denom) : 0.f;
        vectors[nc * K + k] = v;
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)   // the same vector into every peer GPU's gather buffer (NVLink stores)
            if (p < peers.n) peers.dst[p][nc * K + k] = v;
        s += v;
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float tot = 0.f;
        for (int w = 0; w < 8; ++w) tot += part[w];
        vector_sum[nc] = tot;
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
            if (p < peers.n && peers.sum[p] != nullptr) peers.sum[p][nc] = tot;
        valid[nc] = ok ? 1 : 0;
        cnt_all[nc] = ca;
        cnt_gated[nc] = cg;
    }
}

// one CTA per class; images in order; the reference's own fp32 operation sequence.  The prototype row lives
// in registers for the whole walk (R channels per thread, K <= 256 R).  Images are taken S = 16/R at a time:
// all S*R loads of a group are issued together (the update itself is a chain of dependent fp32 operations per
// channel, so latency, not bandwidth, is what has to be hidden), then the group is applied in order from
// registers.  With fin.sums != NULL the kernel first turns segmented sums into mean vectors (proto_finalize
// fused in: counts from the tile records, vector = sum / count, the validity rules), writing the same
// tensors tss_proto_finalize produces.
constexpr int kUpdR = 8;      // max channels per thread: K <= 2048
struct FinalizeArgs {
    const float* sums;
    const TileRecord* records;
    int64_t tiles;
    int min_pixels;
    int32_t* cnt_all;
    int32_t* cnt_gated;
};
template <int R>
__global__ void __launch_bounds__(256)
proto_update_kernel(float* __restrict__ protos, float* __restrict__ num, float* vectors, uint8_t* valid,
                    float* vector_sum, int64_t N, int C, int64_t K, int mode, int start_mean, float m,
                    float one_minus_m, float cap, const FinalizeArgs fin) {
    constexpr int S = 16 / R;
    __shared__ float red[8][S];
    __shared__ int cnt_g[S], cnt_a[S];
    __shared__ uint8_t ok[S];
    const int c = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool fused = fin.sums != nullptr;
    const bool need_sum = fused || vector_sum == nullptr;  // the `vector.sum() == 0` test is evaluated here
    float cnt = num[c];
    float p[R];
#pragma unroll
    for (int r = 0; r < R; ++r) p[r] = (tid + r * 256 < K) ? protos[c * K + tid + r * 256] : 0.f;

    // group g+1's vectors are loaded while group g is applied (the walk over the images is a latency chain)
    const float* src = fused ? fin.sums : vectors;
    float vnext[S][R];
    auto load_group = [&](int64_t n0) {
#pragma unroll
        for (int i = 0; i < S; ++i) {
            const int64_t nc = (n0 + i) * C + c;
#pragma unroll
            for (int r = 0; r < R; ++r) vnext[i][r] = (n0 + i < N && tid + r * 256 < K) ? src[nc * K + tid + r * 256] : 0.f;
        }
    };
    load_group(0);
    for (int64_t n0 = 0; n0 < N; n0 += S) {
        const int len = static_cast<int>(min(static_cast<int64_t>(S), N - n0));
        if (fused) {
            for (int i = warp; i < len; i += 8) {
                int g, a;
                tile_counts(fin.records + (n0 + i) * fin.tiles, fin.tiles, c, lane, g, a);
                if (lane == 0) { cnt_g[i] = g; cnt_a[i] = a; }
            }
            __syncthreads();
        }
        float v[S][R];
#pragma unroll
        for (int i = 0; i < S; ++i)
#pragma unroll
            for (int r = 0; r < R; ++r) v[i][r] = vnext[i][r];
        if (n0 + S < N) load_group(n0 + S);
        if (fused) {
#pragma unroll
            for (int i = 0; i < S; ++i) {
                if (i < len) {
                    const int64_t nc = (n0 + i) * C + c;
                    const bool good = cnt_g[i] > 0 && cnt_a[i] >= fin.min_pixels;
                    const float denom = static_cast<float>(cnt_g[i]);
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        v[i][r] = good ? __fdiv_rn(v[i][r], denom) : 0.f;
                        if (tid + r * 256 < K) vectors[nc * K + tid + r * 256] = v[i][r];
                    }
                }
            }
        }
        if (need_sum) {
#pragma unroll
            for (int i = 0; i < S; ++i) {
                float sacc = 0.f;
#pragma unroll
                for (int r = 0; r < R; ++r) sacc += v[i][r];
                sacc = warp_sum(sacc);
                if (lane == 0) red[warp][i] = sacc;
            }
        }
        __syncthreads();
        if (tid < len) {
            const int64_t nc = (n0 + tid) * C + c;
            float tot = 0.f;
            if (need_sum) {
                for (int w = 0; w < 8; ++w) tot += red[w][tid];
            } else {
                tot = vector_sum[nc];
            }
            bool good;
            if (fused) {
                good = cnt_g[tid] > 0 && cnt_a[tid] >= fin.min_pixels;
                vector_sum[nc] = tot;
                valid[nc] = good ? 1 : 0;
                fin.cnt_all[nc] = cnt_a[tid];
                fin.cnt_gated[nc] = cnt_g[tid];
            } else {
                good = valid == nullptr || valid[nc];
            }
            ok[tid] = (good && tot != 0.f) ? 1 : 0;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < S; ++i) {
            if (i < len && ok[i]) {
                const bool mean = (mode == TSS_UPDATE_MEAN) || (start_mean && cnt < 100.f);
                if (mean) {
                    const float cnt1 = cnt + 1.f;
#pragma unroll
                    for (int r = 0; r < R; ++r) p[r] = __fdiv_rn(__fadd_rn(__fmul_rn(p[r], cnt), v[i][r]), cnt1);
                    cnt = fminf(cnt1, cap);
                } else {
#pragma unroll
                    for (int r = 0; r < R; ++r) p[r] = __fadd_rn(__fmul_rn(p[r], one_minus_m), __fmul_rn(m, v[i][r]));
                    cnt = fminf(cnt + 1.f, cap);
                }
            }
        }
        __syncthreads();  // ok[] / cnt_*[] are rewritten by the next group
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (tid + r * 256 < K) protos[c * K + tid + r * 256] = p[r];
    if (tid == 0) num[c] = cnt;
}

template <typename... A>
static void launch_update(int64_t K, int C, cudaStream_t st, A... args) {
    if (K <= 256) proto_update_kernel<1><<<C, 256, 0, st>>>(args...);
    else if (K <= 512) proto_update_kernel<2><<<C, 256, 0, st>>>(args...);
    else if (K <= 1024) proto_update_kernel<4><<<C, 256, 0, st>>>(args...);
    else proto_update_kernel<8><<<C, 256, 0, st>>>(args...);
}

// closed-form weights of the same recurrence, for the sharded (all-reduce) form
__global__ void __launch_bounds__(256)
proto_update_partial_kernel(const float* __restrict__ vectors, const uint8_t* __restrict__ valid,
                            const float* __restrict__ vector_sum, int64_t N, int C, int64_t K, int mode,
                            int start_mean, double m, double cap, const float* __restrict__ num_before,
                            const int32_t* __restrict__ first_index, const int32_t* __restrict__ total,
                            int32_t* __restrict__ contrib_count, float* __restrict__ partial,
                            float* __restrict__ scale0, float* __restrict__ num_after) {
    extern __shared__ double wts[];  // [N] weight of image n's vector (0 when it does not contribute)
    const int c = blockIdx.x;
    __shared__ int m_local;
    if (threadIdx.x == 0) {
        int cnt = 0;
        for (int64_t n = 0; n < N; ++n) cnt += (valid[n * C + c] && vector_sum[n * C + c] != 0.f) ? 1 : 0;
        m_local = cnt;
        if (contrib_count) contrib_count[c] = cnt;
    }
    __syncthreads();
    if (first_index == nullptr) return;
    if (threadIdx.x == 0) {
        const int first = first_index[c], M = total[c];
        const double n0 = static_cast<double>(num_before[c]);
        // walk the global sequence backwards: suffix = product of a_i for i > j
        double suffix = 1.0;
        int local = m_local - 1;
        // image order of local contributions
        int64_t n_of = N - 1;
        for (int j = M - 1; j >= 0; --j) {
            const double nj = fmin(n0 + j, cap);
            const bool mean = (mode == TSS_UPDATE_MEAN) || (start_mean && nj < 100.0);
            const double a = mean ? nj / (nj + 1.0) : 1.0 - m;
            const double b = mean ? 1.0 / (nj + 1.0) : m;
            if (j >= first && j < first + m_local) {
                while (!(valid[n_of * C + c] && vector_sum[n_of * C + c] != 0.f)) { wts[n_of] = 0.0; --n_of; }
                wts[n_of] = b * suffix;
                --n_of;
                --local;
            }
            suffix *= a;
        }
        for (; n_of >= 0; --n_of) wts[n_of] = 0.0;
        scale0[c] = static_cast<float>(suffix);
        num_after[c] = static_cast<float>(fmin(n0 + M, cap));
    }
    __syncthreads();
    for (int64_t k = threadIdx.x; k < K; k += 256) {
        double s = 0.0;
        for (int64_t n = 0; n < N; ++n) {
            const double w = wts[n];
            if (w != 0.0) s += w * static_cast<double>(vectors[(n * C + c) * K + k]);
        }
        partial[c * K + k] = static_cast<float>(s);
    }
}

}  // namespace tss

extern "C" size_t tss_proto_records_bytes(int64_t n_img, int64_t plane) {
    if (n_img <= 0 || plane <= 0) return 0;
    const int64_t tiles = (plane + tss::kPT - 1) / tss::kPT;
    return static_cast<size_t>(n_img * tiles) * tss::kRecBytes;
}

extern "C" int tss_proto_labels(const float* logits, const float* labels_val, int64_t n_img, int num_classes,
                                int64_t plane, int overflow_channel, int use_conf_gate, float conf_thresh,
                                uint8_t* seg_label, void* records, float* sums_to_zero, int64_t sums_elems,
                                void* stream) {
    using namespace tss;
    TSS_REQUIRE(logits && seg_label && records, TSS_E_ARG, "tss_proto_labels: null pointer");
    TSS_REQUIRE(n_img >= 0 && plane > 0 && n_img <= 65535, TSS_E_ARG, "tss_proto_labels: bad extents");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES, "tss_proto_labels: num_classes %d", num_classes);
    TSS_REQUIRE((reinterpret_cast<uintptr_t>(records) & 15) == 0, TSS_E_ALIGN, "tss_proto_labels: records must be 16-byte aligned");
    TSS_REQUIRE(sums_elems >= 0 && (sums_elems == 0 || sums_to_zero != nullptr), TSS_E_ARG, "tss_proto_labels: sums_elems without buffer");
    if (n_img == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    const int64_t tiles = (plane + kPT - 1) / kPT;
    dim3 grid(static_cast<unsigned>(tiles), static_cast<unsigned>(n_img));
    auto* rec = static_cast<TileRecord*>(records);
    // the 228 KB shared-memory split of the accumulate kernel that follows: this kernel then runs beside it (and beside a
    // co-resident pseudo-label kernel) on the same SMs instead of waiting for a reconfiguration
    cudaFuncSetAttribute(proto_label_kernel<13>, cudaFuncAttributePreferredSharedMemoryCarveout, static_cast<int>(cudaSharedmemCarveoutMaxShared));
    cudaFuncSetAttribute(proto_label_kernel<19>, cudaFuncAttributePreferredSharedMemoryCarveout, static_cast<int>(cudaSharedmemCarveoutMaxShared));
    cudaFuncSetAttribute(proto_label_kernel<TSS_MAX_CLASSES>, cudaFuncAttributePreferredSharedMemoryCarveout, static_cast<int>(cudaSharedmemCarveoutMaxShared));
    if (num_classes <= 13)
        proto_label_kernel<13><<<grid, 256, 0, st>>>(logits, labels_val, num_classes, plane, overflow_channel, use_conf_gate, conf_thresh, seg_label, rec, sums_to_zero, sums_elems);
    else if (num_classes <= 19)
        proto_label_kernel<19><<<grid, 256, 0, st>>>(logits, labels_val, num_classes, plane, overflow_channel, use_conf_gate, conf_thresh, seg_label, rec, sums_to_zero, sums_elems);
    else
        proto_label_kernel<TSS_MAX_CLASSES><<<grid, 256, 0, st>>>(logits, labels_val, num_classes, plane, overflow_channel, use_conf_gate, conf_thresh, seg_label, rec, sums_to_zero, sums_elems);
    TSS_LAUNCH_CHECK("tss_proto_labels");
    return TSS_OK;
}

namespace tss {
template <int CT, int NC>
static int launch_accum(const float* feat, const TileRecord* rec, int64_t N, int64_t K, int64_t P, int C, float* sums,
                        cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(proto_accum_kernel<CT, NC>, cudaFuncAttributeMaxDynamicSharedMemorySize, AccSmem::total);
    if (e != cudaSuccess) return cuda_status(e, "tss_proto_accumulate: smem attribute");
    const int64_t jobs = N * (K / kKT) * ((P + kPT - 1) / kPT);
    const int grid = static_cast<int>(jobs < kNumSMs ? jobs : kNumSMs);
    proto_accum_kernel<CT, NC><<<grid, AccCfg<NC>::kThreads, AccSmem::total, st>>>(feat, rec, N, K, P, C, sums);
    TSS_LAUNCH_CHECK("tss_proto_accumulate");
    return TSS_OK;
}
template <int CT>
static int launch_accum_nc(int nc, const float* feat, const TileRecord* rec, int64_t N, int64_t K, int64_t P, int C, float* sums,
                           cudaStream_t st) {
    switch (nc) {
        case 4: return launch_accum<CT, 4>(feat, rec, N, K, P, C, sums, st);
        case 6: return launch_accum<CT, 6>(feat, rec, N, K, P, C, sums, st);
        case 8: return launch_accum<CT, 8>(feat, rec, N, K, P, C, sums, st);
        default: return launch_accum<CT, 16>(feat, rec, N, K, P, C, sums, st);
    }
}
}  // namespace tss

extern "C" int tss_proto_accumulate(const float* feat, const void* records, int64_t n_img, int64_t channels,
                                    int64_t plane, int num_classes, float* sums, int zero_sums, int consumer_warps,
                                    void* stream) {
    using namespace tss;
    TSS_REQUIRE(feat && records && sums, TSS_E_ARG, "tss_proto_accumulate: null pointer");
    TSS_REQUIRE(n_img >= 0 && plane > 0 && channels > 0, TSS_E_ARG, "tss_proto_accumulate: bad extents");
    TSS_REQUIRE(channels % kKT == 0, TSS_E_ARG, "tss_proto_accumulate: channels (%lld) must be a multiple of 32", static_cast<long long>(channels));
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES, "tss_proto_accumulate: num_classes %d", num_classes);
    TSS_REQUIRE((reinterpret_cast<uintptr_t>(feat) & 15) == 0 && (reinterpret_cast<uintptr_t>(records) & 15) == 0, TSS_E_ALIGN,
                "tss_proto_accumulate: feat and records must be 16-byte aligned");
    TSS_REQUIRE(consumer_warps == 0 || consumer_warps == 4 || consumer_warps == 6 || consumer_warps == 8 || consumer_warps == 16, TSS_E_ARG,
                "tss_proto_accumulate: consumer_warps must be 0 (default 16), 4, 6, 8 or 16");
    if (n_img == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    if (zero_sums) {
        cudaError_t e = cudaMemsetAsync(sums, 0, sizeof(float) * n_img * num_classes * channels, st);
        if (e != cudaSuccess) return cuda_status(e, "tss_proto_accumulate: memset");
    }
    auto* rec = static_cast<const TileRecord*>(records);
    if (num_classes <= 13) return launch_accum_nc<13>(consumer_warps, feat, rec, n_img, channels, plane, num_classes, sums, st);
    if (num_classes <= 19) return launch_accum_nc<19>(consumer_warps, feat, rec, n_img, channels, plane, num_classes, sums, st);
    return launch_accum_nc<TSS_MAX_CLASSES>(consumer_warps, feat, rec, n_img, channels, plane, num_classes, sums, st);
}

extern "C" int tss_proto_finalize(const float* sums, const void* records, int64_t n_img, int num_classes,
                                  int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                                  int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum, void* stream) {
    using namespace tss;
    TSS_REQUIRE(sums && records && count_all && count_gated && vectors && valid && vector_sum, TSS_E_ARG, "tss_proto_finalize: null pointer");
    TSS_REQUIRE(n_img >= 0 && channels > 0 && plane > 0 && num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_ARG, "tss_proto_finalize: bad extents");
    if (n_img == 0) return TSS_OK;
    proto_finalize_kernel<<<static_cast<unsigned>(n_img * num_classes), 256, 0, as_stream(stream)>>>(
        sums, static_cast<const TileRecord*>(records), (plane + kPT - 1) / kPT, num_classes, channels, min_pixels, count_all,
        count_gated, vectors, valid, vector_sum, PeerList{{nullptr}, {nullptr}, 0});
    TSS_LAUNCH_CHECK("tss_proto_finalize");
    return TSS_OK;
}

extern "C" int tss_proto_finalize_scatter(const float* sums, const void* records, int64_t n_img, int num_classes,
                                          int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                                          int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum,
                                          float* const* peer_vectors, float* const* peer_vector_sums, int n_peers,
                                          void* stream) {
    using namespace tss;
    TSS_REQUIRE(sums && records && count_all && count_gated && vectors && valid && vector_sum, TSS_E_ARG, "tss_proto_finalize_scatter: null pointer");
    TSS_REQUIRE(n_img >= 0 && channels > 0 && plane > 0 && num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_ARG, "tss_proto_finalize_scatter: bad extents");
    TSS_REQUIRE(n_peers >= 0 && n_peers <= kMaxPeers && (n_peers == 0 || peer_vectors), TSS_E_ARG, "tss_proto_finalize_scatter: n_peers %d outside [0, %d]", n_peers, kMaxPeers);
    if (n_img == 0) return TSS_OK;
    PeerList pl{{nullptr}, {nullptr}, n_peers};
    for (int p = 0; p < n_peers; ++p) {
        TSS_REQUIRE(peer_vectors[p] != nullptr, TSS_E_ARG, "tss_proto_finalize_scatter: null peer pointer %d", p);
        pl.dst[p] = peer_vectors[p];
        pl.sum[p] = peer_vector_sums ? peer_vector_sums[p] : nullptr;
    }
    proto_finalize_kernel<<<static_cast<unsigned>(n_img * num_classes), 256, 0, as_stream(stream)>>>(
        sums, static_cast<const TileRecord*>(records), (plane + kPT - 1) / kPT, num_classes, channels, min_pixels, count_all,
        count_gated, vectors, valid, vector_sum, pl);
    TSS_LAUNCH_CHECK("tss_proto_finalize_scatter");
    return TSS_OK;
}

namespace tss {
static int check_update_args(const char* who, const void* protos, const void* num, const void* vectors, int64_t n_img,
                             int num_classes, int64_t channels, int mode) {
    TSS_REQUIRE(protos && num && vectors, TSS_E_ARG, "%s: null pointer", who);
    TSS_REQUIRE(n_img >= 0 && channels > 0 && num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_ARG, "%s: bad extents", who);
    TSS_REQUIRE(mode == TSS_UPDATE_MEAN || mode == TSS_UPDATE_MOVING_AVERAGE, TSS_E_ARG, "%s: bad mode %d", who, mode);
    TSS_REQUIRE(channels <= 256 * kUpdR, TSS_E_ARG, "%s: at most %d channels", who, 256 * kUpdR);
    return TSS_OK;
}
}  // namespace tss

extern "C" int tss_proto_update(float* protos, float* num, const float* vectors, const uint8_t* valid,
                                const float* vector_sum, int64_t n_img, int num_classes, int64_t channels, int mode,
                                int start_mean, double momentum, double cap, void* stream) {
    using namespace tss;
    const int rc = check_update_args("tss_proto_update", protos, num, vectors, n_img, num_classes, channels, mode);
    if (rc != TSS_OK) return rc;
    if (n_img == 0) return TSS_OK;
    // python evaluates (1 - momentum) in double before the fp32 tensor multiply
    launch_update(channels, num_classes, as_stream(stream),
        protos, num, const_cast<float*>(vectors), const_cast<uint8_t*>(valid), const_cast<float*>(vector_sum), n_img,
        num_classes, channels, mode, start_mean, static_cast<float>(momentum), static_cast<float>(1.0 - momentum),
        static_cast<float>(cap), FinalizeArgs{nullptr, nullptr, 0, 0, nullptr, nullptr});
    TSS_LAUNCH_CHECK("tss_proto_update");
    return TSS_OK;
}

extern "C" int tss_proto_finalize_update(const float* sums, const void* records, int64_t n_img, int num_classes,
                                         int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                                         int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum,
                                         float* protos, float* num, int mode, int start_mean, double momentum,
                                         double cap, void* stream) {
    using namespace tss;
    TSS_REQUIRE(sums && records && count_all && count_gated && valid && vector_sum && plane > 0, TSS_E_ARG,
                "tss_proto_finalize_update: null pointer");
    const int rc = check_update_args("tss_proto_finalize_update", protos, num, vectors, n_img, num_classes, channels, mode);
    if (rc != TSS_OK) return rc;
    if (n_img == 0) return TSS_OK;
    launch_update(channels, num_classes, as_stream(stream),
        protos, num, vectors, valid, vector_sum, n_img, num_classes, channels, mode, start_mean, static_cast<float>(momentum),
        static_cast<float>(1.0 - momentum), static_cast<float>(cap),
        FinalizeArgs{sums, static_cast<const TileRecord*>(records), (plane + kPT - 1) / kPT, min_pixels, count_all, count_gated});
    TSS_LAUNCH_CHECK("tss_proto_finalize_update");
    return TSS_OK;
}

extern "C" int tss_proto_update_partial(const float* vectors, const uint8_t* valid, const float* vector_sum,
                                        int64_t n_img, int num_classes, int64_t channels, int mode, int start_mean,
                                        double momentum, double cap, const float* num_before,
                                        const int32_t* first_index, const int32_t* total, int32_t* contrib_count,
                                        float* partial, float* scale0, float* num_after, void* stream) {
    using namespace tss;
    TSS_REQUIRE(vectors && valid && vector_sum, TSS_E_ARG, "tss_proto_update_partial: null pointer");
    TSS_REQUIRE(n_img >= 0 && n_img <= 16384 && channels > 0 && num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_ARG,
                "tss_proto_update_partial: bad extents (at most 16384 images per call)");
    TSS_REQUIRE(first_index != nullptr || contrib_count != nullptr, TSS_E_ARG, "tss_proto_update_partial: nothing to do");
    TSS_REQUIRE(first_index == nullptr || (total && partial && scale0 && num_after && num_before), TSS_E_ARG,
                "tss_proto_update_partial: weighted pass needs total, partial, scale0, num_before, num_after");
    const size_t sh = sizeof(double) * static_cast<size_t>(n_img > 0 ? n_img : 1);
    cudaError_t e = cudaFuncSetAttribute(proto_update_partial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    if (e != cudaSuccess) return cuda_status(e, "tss_proto_update_partial: smem attribute");
    proto_update_partial_kernel<<<num_classes, 256, sh, as_stream(stream)>>>(vectors, valid, vector_sum, n_img, num_classes, channels,
                                                                             mode, start_mean, momentum, cap, num_before, first_index,
                                                                             total, contrib_count, partial, scale0, num_after);
    TSS_LAUNCH_CHECK("tss_proto_update_partial");
    return TSS_OK;
}

// ------------------------------------------------------------------------------------------------
// tss_class_means_noalloc: the three steps of A5 (labels -> segmented sums -> means) for callers that cannot allocate.
// No counterpart in the reference.  The scratch the three kernels hand to
// each other (tile records, per-pixel segment labels, the [N,C,K] sums) lives in ONE file-scope device array, so the
// call performs no cudaMalloc / cudaFree and takes no caller workspace.  In host code `g_noalloc_scratch` is only a
// stand-in for the symbol: its device address is looked up once per device with cudaGetSymbolAddress (with that device
// current) and that pointer is what the kernels receive.
namespace tss {
static __device__ __align__(256) unsigned char g_noalloc_scratch[TSS_NOALLOC_SCRATCH_BYTES];

constexpr int kNoallocMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static unsigned char* g_noalloc_ptr[kNoallocMaxDevices] = {};

static int noalloc_scratch(unsigned char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_status(e, "tss_class_means_noalloc: cudaGetDevice");
    TSS_REQUIRE(dev >= 0 && dev < kNoallocMaxDevices, TSS_E_ARG, "tss_class_means_noalloc: device ordinal %d outside [0, %d)", dev,
                kNoallocMaxDevices);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        unsigned char* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);   // this device's copy of the array
        if (e != cudaSuccess) return cuda_status(e, "tss_class_means_noalloc: cudaGetSymbolAddress");
        TSS_REQUIRE(p != nullptr, TSS_E_ARG, "tss_class_means_noalloc: null scratch address");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return TSS_OK;
}

static inline size_t up256(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }
}  // namespace tss

extern "C" int tss_class_means_noalloc(const float* feat, const float* logits, int64_t n_img, int num_classes,
                                       int64_t channels, int64_t plane, int min_pixels, int32_t* count_all,
                                       int32_t* count_gated, float* vectors, uint8_t* valid, float* vector_sum,
                                       void* stream) {
    using namespace tss;
    TSS_REQUIRE(feat && logits && count_all && count_gated && vectors && valid && vector_sum, TSS_E_ARG,
                "tss_class_means_noalloc: null pointer");
    TSS_REQUIRE(n_img >= 0 && n_img <= 65535 && plane > 0 && channels > 0 && channels % 32 == 0, TSS_E_ARG,
                "tss_class_means_noalloc: bad extents (channels must be a multiple of 32)");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES, "tss_class_means_noalloc: num_classes %d",
                num_classes);
    if (n_img == 0) return TSS_OK;
    // scratch layout: [records | segment labels | sums], each part rounded up to 256 bytes
    const size_t rec_bytes = up256(tss_proto_records_bytes(n_img, plane));
    const size_t seg_bytes = up256(static_cast<size_t>(n_img) * static_cast<size_t>(plane));
    const size_t sums_elems = static_cast<size_t>(n_img) * num_classes * static_cast<size_t>(channels);
    const size_t need = rec_bytes + seg_bytes + up256(sums_elems * sizeof(float));
    TSS_REQUIRE(need <= static_cast<size_t>(TSS_NOALLOC_SCRATCH_BYTES), TSS_E_ARG,
                "tss_class_means_noalloc: this batch needs %zu bytes of scratch, the built-in array holds %zu "
                "(TSS_NOALLOC_SCRATCH_BYTES): use tss_proto_labels / _accumulate / _finalize with caller memory",
                need, static_cast<size_t>(TSS_NOALLOC_SCRATCH_BYTES));
    unsigned char* base = nullptr;
    if (int r = noalloc_scratch(&base)) return r;
    void* records = base;
    uint8_t* seg = base + rec_bytes;
    float* sums = reinterpret_cast<float*>(base + rec_bytes + seg_bytes);
    // the array keeps the previous call's contents: the label kernel rewrites every record and segment label it will be
    // read for and zeroes the sums in the same launch (sums_to_zero), so nothing else has to be cleared
    if (int r = tss_proto_labels(logits, nullptr, n_img, num_classes, plane, num_classes, 0, -1.0f, seg, records, sums,
                                 static_cast<int64_t>(sums_elems), stream)) return r;
    if (int r = tss_proto_accumulate(feat, records, n_img, channels, plane, num_classes, sums, 0, 0, stream)) return r;
    return tss_proto_finalize(sums, records, n_img, num_classes, channels, plane, min_pixels, count_all, count_gated, vectors,
                              valid, vector_sum, stream);
}
