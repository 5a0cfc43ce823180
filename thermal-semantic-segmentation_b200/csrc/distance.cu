// A7/A8/A10: feature-to-prototype L2 distance, soft weights, label rectification
// (self_training.py:160-165, :139-142, :106-112).
//
// The reference loops over classes and re-reads the whole [N,K,H,W] feature tensor C times; here feat
// is read ONCE.  The contraction is [px, K] x [K, C <= 20]: with so few classes it is bound by HBM and
// the fp32 FMA pipe, not a tensor-core shape (BASELINE.json north_star), so it runs on CUDA cores in the
// reference's direct-difference form (p - f)^2 with packed fp32x2 math (FADD2/FFMA2: two channels per
// issue slot), prototypes broadcast from shared memory as 128-bit loads.
//
// CTA = 8 warps, persistent (one per SM, up to 255 registers per thread).  The whole prototype matrix is staged
// ONCE per CTA when it fits shared memory (156 KB at C = 19, K = 2048).  A lane owns FOUR pixels (p, p+32, p+64,
// p+96 of a 128-pixel warp tile) packed as two f32x2 pairs, and keeps their 2 x C squared-distance accumulators in
// registers; a prototype value is a scalar operand broadcast to both halves of FADD2 / FFMA2.  One 128-bit
// broadcast shared load (4 channels of one class) then feeds 16 packed FP32 instructions: the shared-memory
// return path (128 B/clk/SM) stays at half the FP32 pipe's time -- with one or two pixels per lane it is the
// bound.  K is split S ways across warps (S = 1 when the batch has enough tiles to fill the machine), the S
// partials meet in shared memory.  Feature loads run D four-channel groups ahead of the arithmetic in a rotating
// register pipeline: the kernel is FP32-pipe bound (2 flop-slots per pixel, channel and class), what has to be
// hidden is the HBM latency.  sqrt / min / softmax / rectify finish per pixel straight from registers.
#include "common.cuh"

namespace tss {

constexpr int kDThreads = 256;
constexpr int kDWarps = 8;
constexpr int kDDepth = 4;    // four-channel groups in flight per warp (16 loads each)
constexpr int kDWarpPx = 128; // pixels per warp tile

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

// per-pixel epilogue: weights (A8) and rectified label (A10) from the C distances d[] of pixel p of image n
template <int CT>
__device__ __forceinline__ void distance_epilogue(const DistArgs& a, float (&d)[CT], int64_t n, int64_t p) {
    float dmin = INFINITY;
#pragma unroll
    for (int c = 0; c < CT; ++c)
        if (c < a.C) dmin = fminf(dmin, d[c]);
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

// Persistent CTAs.  Warp w: pixel group w / S (128 pixels), K part w % S.  When the whole prototype matrix fits
// shared memory (KC == K) it is staged once per CTA; otherwise K is processed in chunks of KC channels that are
// re-staged per tile.
template <int CT>
__global__ void __launch_bounds__(kDThreads, 1) distance_kernel(const DistArgs a, int64_t tiles, int S) {
    extern __shared__ __align__(16) float sm[];
    const int KC = a.KC, PS = KC + 4;  // prototype row stride in shared memory (16-byte aligned rows)
    float* ps = sm;                    // [CT][PS]
    float* red = sm + CT * PS;         // [8][CT][2][32] split-K partials of one pixel pair (S > 1)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = kDWarps / S;         // pixel groups per CTA
    const int grp = warp / S, part = warp - grp * S;
    const int tile_px = kDWarpPx * G;
    const int64_t tiles_per_img = (a.P + tile_px - 1) / tile_px;
    const bool resident = KC == a.K;
    const int KW = KC / S;             // channels per warp per chunk (multiple of 4: K % 32 == 0, S <= 8)
    const int kb = part * KW;

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
        const int64_t p0 = (tile - n * tiles_per_img) * tile_px + grp * kDWarpPx + lane;
        const float* fimg = a.feat + n * a.K * a.P;
        int64_t po[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) po[j] = min(p0 + 32 * j, a.P - 1);
        f32x2 acc[CT][2];
#pragma unroll
        for (int c = 0; c < CT; ++c) { acc[c][0] = 0ull; acc[c][1] = 0ull; }

        for (int64_t kc0 = 0; kc0 < a.K; kc0 += KC) {
            if (!resident) {
                __syncthreads();
                stage(kc0);
                __syncthreads();
            }
            const float* f = fimg + (kc0 + kb) * a.P;
            const float* pk = ps + kb;
            // raw loaded values [group][channel][pixel]; they are paired into f32x2 operands only where they are
            // consumed -- a pack (MOV) next to the load would make the in-order warp wait for the load at once
            float r[kDDepth][4][4];
            auto load_group = [&](float (&dst)[4][4], int k0) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float* row = f + (k0 + q) * a.P;
#pragma unroll
                    for (int j4 = 0; j4 < 4; ++j4) dst[q][j4] = ldg_stream_f(row + po[j4]);
                }
            };
#pragma unroll
            for (int j = 0; j < kDDepth; ++j)
                if (4 * j < KW) load_group(r[j], 4 * j);
            for (int kk = 0; kk < KW; kk += 4 * kDDepth) {
#pragma unroll
                for (int j = 0; j < kDDepth; ++j) {
                    const int k0 = kk + 4 * j;
                    if (k0 < KW) {  // warp-uniform
                        f32x2 fp[4][2];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            fp[q][0] = pack2(r[j][q][0], r[j][q][1]);
                            fp[q][1] = pack2(r[j][q][2], r[j][q][3]);
                        }
#pragma unroll
                        for (int c = 0; c < CT; ++c) {
                            const float4 pr = *reinterpret_cast<const float4*>(pk + c * PS + k0);
                            const float pv[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const f32x2 pp = pack2(pv[q], pv[q]);  // scalar operand, broadcast to both halves
                                f32x2 d;
                                d = sub2(pp, fp[q][0]); acc[c][0] = fma2(d, d, acc[c][0]);
                                d = sub2(pp, fp[q][1]); acc[c][1] = fma2(d, d, acc[c][1]);
                            }
                        }
                        const int kn = k0 + 4 * kDDepth;
                        if (kn < KW) load_group(r[j], kn);
                    }
                }
            }
        }
        // squared distances of the lane's four pixels
        float d2[4][CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            unpack2(acc[c][0], d2[0][c], d2[1][c]);
            unpack2(acc[c][1], d2[2][c], d2[3][c]);
        }
        if (S > 1) {  // CTA-uniform: combine the S partials, one pixel pair per round (shared-memory budget)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                __syncthreads();  // the previous round's partials have been consumed
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                    red[((warp * CT + c) * 2 + 0) * 32 + lane] = d2[2 * h][c];
                    red[((warp * CT + c) * 2 + 1) * 32 + lane] = d2[2 * h + 1][c];
                }
                __syncthreads();
                if (part == 0) {
                    for (int s2 = 1; s2 < S; ++s2) {
#pragma unroll
                        for (int c = 0; c < CT; ++c) {
                            d2[2 * h][c] += red[(((warp + s2) * CT + c) * 2 + 0) * 32 + lane];
                            d2[2 * h + 1][c] += red[(((warp + s2) * CT + c) * 2 + 1) * 32 + lane];
                        }
                    }
                }
            }
        }
        if (part == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t p = p0 + 32 * j;
                if (p < a.P) {
                    float d[CT];
#pragma unroll
                    for (int c = 0; c < CT; ++c) d[c] = sqrtf(d2[j][c]);
                    distance_epilogue<CT>(a, d, n, p);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Thread-level-parallel form.  ONE large CTA per SM shares the whole prototype matrix in shared memory and every
// warp works on its own pixel unit with no CTA-wide synchronisation after the prologue.  A lane owns PX (2 or 4)
// pixels as packed pairs, so a thread needs CT*PX accumulator registers + one 4-channel group of loaded values
// and the SM holds 12-32 warps: the latency of the feature loads is hidden by warp switching instead of by a
// register pipeline (whose loads share the warp's few scoreboards, so waiting for the oldest group also waits
// for the newest).  Load balance comes from an INTRA-warp K split: the warp's lanes form SI (1, 2 or 4) parts
// of L = 32/SI lanes; all parts cover the same PX*L pixels, part s reads channels [s*K/SI, (s+1)*K/SI), and the
// partial sums meet in log2(SI) xor-shuffles -- no shared-memory exchange, no barrier, and units small enough
// (down to 16 pixels) that the last round of the persistent loop is nearly full.  A quarter warp (8 lanes) always
// lies inside one part, so the float4 prototype read stays a broadcast.
// PF: the next 4-channel group is loaded into a second register set while the current one is used.  It costs
// 4 PX registers (fewer warps per SM) and buys 7-15 % per pixel-channel; which of the two wins depends on how well
// the resulting number of warp slots divides the problem (tlp_cost).
template <int CT, int PX, int PF>
struct TlpCfg {
    static constexpr int kAcc = CT * PX + PF * 4 * PX;
    static constexpr int kWarps = kAcc <= 26 ? 32 : kAcc <= 40 ? 24 : kAcc <= 56 ? 20 : kAcc <= 80 ? 16 : 12;
    static constexpr int kThreads = kWarps * 32;
};

// EXP (expanded form): the loop accumulates f.p (ONE packed FMA per pixel pair, channel and class instead of a
// subtract + FMA) and |f|^2 once per pixel; d^2 = |f|^2 + |p|^2 - 2 f.p.  The expansion cancels when a feature
// lies near a prototype, so every (pixel, class) whose d^2 comes out below kExpTheta (|f|^2 + |p|^2) is
// recomputed in the reference's direct-difference form by the whole warp (exact_d2_pixel): the fp32
// accumulation error of the three sums (about 1e-6 relative to |f|^2 + |p|^2, measured; 4e-6 at four sigma) then
// stays below 1e-5 relative to d^2 everywhere else.
constexpr float kExpTheta = 0.5f;
#ifndef TSS_STREAM_AUTO_BIG
#define TSS_STREAM_AUTO_BIG 3   // tile shape of the streaming kernel for large batches (profiles/r02_distance_forms.txt)
#endif

template <int CT>
__device__ __noinline__ void exact_d2_pixel(const DistArgs& a, int64_t n, int64_t p, unsigned mask, float (&out)[CT]) {
    const int lane = threadIdx.x & 31;
    float s[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) s[c] = 0.f;
    const float* f = a.feat + n * a.K * a.P + p;
    for (int64_t k = lane; k < a.K; k += 32) {   // lane = channel (mod 32): 32 rows of the pixel per step
        const float x = __ldg(f + k * a.P);
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if ((mask >> c) & 1u) {
                const float d = __fsub_rn(__ldg(a.protos + c * a.K + k), x);
                s[c] = fmaf(d, d, s[c]);
            }
    }
#pragma unroll
    for (int c = 0; c < CT; ++c) out[c] = ((mask >> c) & 1u) ? warp_sum(s[c]) : 0.f;
}

template <int CT, int PX, int PF, bool EXP>
__global__ void __launch_bounds__(TlpCfg<CT, PX, PF>::kThreads, 1) distance_tlp_kernel(const DistArgs a, int64_t units, int SI) {
    constexpr int W = TlpCfg<CT, PX, PF>::kWarps;
    constexpr int PP = PX / 2;
    // [CT][K + 4 SI], rows >= C zero.  The K/SI channels of part s start 16 bytes (4 s floats) later than their
    // natural place: the SI addresses of one float4 read (same class, same step, channel offsets s K/SI -- multiples
    // of 128 bytes) would otherwise fall into the same four banks and cost SI wavefronts instead of one.
    extern __shared__ __align__(16) float ps[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int K = static_cast<int>(a.K);
    const int KW = K / SI;
    const int PS = K + 4 * SI;
    for (int i = tid; i < CT * K; i += TlpCfg<CT, PX, PF>::kThreads) {
        const int c = i / K, k = i - c * K;
        ps[c * PS + k + 4 * (k / KW)] = (c < a.C) ? a.protos[i] : 0.f;
    }
    float* pn = ps + CT * PS;   // [CT] |p_c|^2 (EXP), summed in fp64
    if (EXP) {
        for (int c = warp; c < CT; c += W) {
            double sq = 0.0;
            if (c < a.C)
                for (int k = lane; k < K; k += 32) {
                    const double v = static_cast<double>(a.protos[c * K + k]);
                    sq += v * v;
                }
            sq = warp_sum_d(sq);
            if (lane == 0) pn[c] = static_cast<float>(sq);
        }
    }
    __syncthreads();

    const int L = 32 / SI;
    const int part = lane / L, j = lane - part * L;
    const int unit_px = PX * L;
    const int64_t units_per_img = (a.P + unit_px - 1) / unit_px;
    const float* pk = ps + part * (KW + 4);
    const int64_t P = a.P;

    for (int64_t u = static_cast<int64_t>(blockIdx.x) * W + warp; u < units; u += static_cast<int64_t>(gridDim.x) * W) {
        const int64_t n = u / units_per_img;
        const int64_t base = (u - n * units_per_img) * unit_px + j;
        const int64_t first = min(base, P - 1);
        int off[PX];
#pragma unroll
        for (int i = 0; i < PX; ++i) off[i] = static_cast<int>(min(base + i * L, P - 1) - first);
        const float* row = a.feat + (n * a.K + static_cast<int64_t>(part) * KW) * P + first;
        f32x2 acc[CT][PP];
#pragma unroll
        for (int c = 0; c < CT; ++c)
#pragma unroll
            for (int h = 0; h < PP; ++h) acc[c][h] = 0ull;
        f32x2 nrm[2][PP];   // EXP: |f|^2 in two interleaved chains per pixel pair
#pragma unroll
        for (int h = 0; h < PP; ++h) { nrm[0][h] = 0ull; nrm[1][h] = 0ull; }

        // PF = prefetch depth in 4-channel groups (0: loads of a group are issued when it is used)
        constexpr int DS = PF > 0 ? PF : 1;
        float xn[DS][PF > 0 ? 4 : 1][PF > 0 ? PX : 1];
        const float* ld = row;                                              // next group to fetch
        const float* last = row + static_cast<int64_t>(KW - 4) * P;         // the part's last group (always in range)
        if (PF > 0) {
#pragma unroll
            for (int d = 0; d < DS; ++d) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
#pragma unroll
                    for (int i = 0; i < PX; ++i) xn[d][PF > 0 ? q : 0][PF > 0 ? i : 0] = ldg_stream_f(ld + off[i]);
                    ld += P;
                }
            }
        }
#pragma unroll(PF > 0 ? 1 : (PX == 2 ? 2 : 1))
        for (int kk0 = 0; kk0 < KW; kk0 += 4 * DS) {
#pragma unroll
          for (int d = 0; d < DS; ++d) {
            const int kk = kk0 + 4 * d;
            float x[4][PX];
            if (PF > 0) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int i = 0; i < PX; ++i) x[q][i] = xn[d][PF > 0 ? q : 0][PF > 0 ? i : 0];
                // refill this stage with group kk/4 + PF; past the end the last group is re-read (in range, unused)
                const bool inr = kk + 4 * DS < KW;   // warp-uniform
                const float* src = inr ? ld : last;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
#pragma unroll
                    for (int i = 0; i < PX; ++i) xn[d][PF > 0 ? q : 0][PF > 0 ? i : 0] = ldg_stream_f(src + off[i]);
                    src += P;
                }
                if (inr) ld = src;
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
#pragma unroll
                    for (int i = 0; i < PX; ++i) x[q][i] = ldg_stream_f(ld + off[i]);
                    ld += P;
                }
            }
            f32x2 xp[4][PP];
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int h = 0; h < PP; ++h) xp[q][h] = pack2(x[q][2 * h], x[q][2 * h + 1]);
            if (EXP) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
#pragma unroll
                    for (int h = 0; h < PP; ++h) nrm[q & 1][h] = fma2(xp[q][h], xp[q][h], nrm[q & 1][h]);
            }
#pragma unroll
            for (int c = 0; c < CT; ++c) {
                const float4 pr = *reinterpret_cast<const float4*>(pk + c * PS + kk);
                const float pv[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const f32x2 pb = pack2(pv[q], pv[q]);
#pragma unroll
                    for (int h = 0; h < PP; ++h) {
                        if (EXP) {
                            acc[c][h] = fma2(pb, xp[q][h], acc[c][h]);
                        } else {
                            const f32x2 df = sub2(pb, xp[q][h]);
                            acc[c][h] = fma2(df, df, acc[c][h]);
                        }
                    }
                }
            }
          }
        }
        if (EXP) {
#pragma unroll
            for (int h = 0; h < PP; ++h) nrm[0][h] = add2(nrm[0][h], nrm[1][h]);
        }
        for (int o = 16; o >= L; o >>= 1) {  // partial sums of the SI parts (no iteration when SI == 1)
#pragma unroll
            for (int c = 0; c < CT; ++c)
#pragma unroll
                for (int h = 0; h < PP; ++h) acc[c][h] = add2(acc[c][h], __shfl_xor_sync(0xffffffffu, acc[c][h], o));
            if (EXP) {
#pragma unroll
                for (int h = 0; h < PP; ++h) nrm[0][h] = add2(nrm[0][h], __shfl_xor_sync(0xffffffffu, nrm[0][h], o));
            }
        }
        if (!EXP) {
#pragma unroll
            for (int i = 0; i < PX; ++i) {   // pixel i of the unit is finished by part i mod SI
                const int64_t p = base + i * L;
                if ((i & (SI - 1)) == part && p < P) {
                    float d[CT];
#pragma unroll
                    for (int c = 0; c < CT; ++c) {
                        float lo, hi;
                        unpack2(acc[c][i >> 1], lo, hi);
                        d[c] = sqrtf((i & 1) ? hi : lo);
                    }
                    distance_epilogue<CT>(a, d, n, p);
                }
            }
        } else {
            // d^2 = |f|^2 + |p|^2 - 2 f.p per finishing pixel; pairs inside the cancellation region are flagged
            float d2[PX][CT];
            unsigned fl[PX];
            unsigned any = 0;
#pragma unroll
            for (int i = 0; i < PX; ++i) {
                const int64_t p = base + i * L;
                const bool mine = (i & (SI - 1)) == part && p < P;
                float flo, fhi;
                unpack2(nrm[0][i >> 1], flo, fhi);
                const float F = (i & 1) ? fhi : flo;
                fl[i] = 0;
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                    float lo, hi;
                    unpack2(acc[c][i >> 1], lo, hi);
                    const float scale = F + pn[c];
                    d2[i][c] = fmaf(-2.0f, (i & 1) ? hi : lo, scale);
                    if (mine && c < a.C && d2[i][c] < kExpTheta * scale) fl[i] |= 1u << c;
                }
                any |= fl[i];
            }
            if (__any_sync(0xffffffffu, any != 0)) {   // rare: the warp recomputes the flagged pairs exactly
#pragma unroll
                for (int i = 0; i < PX; ++i) {
                    unsigned pend = __ballot_sync(0xffffffffu, fl[i] != 0);
                    while (pend) {
                        const int src = __ffs(pend) - 1;
                        pend &= pend - 1;
                        const unsigned m = __shfl_sync(0xffffffffu, fl[i], src);
                        const int pp = __shfl_sync(0xffffffffu, static_cast<int>(base + i * L), src);
                        float ex[CT];
                        exact_d2_pixel<CT>(a, n, pp, m, ex);
                        if (lane == src) {
#pragma unroll
                            for (int c = 0; c < CT; ++c)
                                if ((m >> c) & 1u) d2[i][c] = ex[c];
                        }
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < PX; ++i) {
                const int64_t p = base + i * L;
                if ((i & (SI - 1)) == part && p < P) {
                    float d[CT];
#pragma unroll
                    for (int c = 0; c < CT; ++c) d[c] = sqrtf(fmaxf(d2[i][c], 0.f));
                    distance_epilogue<CT>(a, d, n, p);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Streaming form (the default for K >= 512; shorter K loops do not amortise the per-unit pipeline fill and epilogue).  The thread-level-parallel kernels above are bound by the latency of their
// own feature loads (ncu: 7 long-scoreboard stall cycles per issued instruction, DRAM 23 %, FMA pipe 35 %): with C
// accumulator registers per pixel a warp can keep only a few KB in flight.  Here the loads are the TMA engine's:
// persistent CTAs, one producer warp that streams (32 channels x 512 pixels) tiles of feat -- and the matching
// [C, 32] slice of the prototypes -- into a 3-stage shared-memory ring with 1-D bulk copies completing on mbarriers
// (200 KB in flight per SM), and eight consumer warps that own 64 pixels each (a lane owns pixels l and l + 32 as one
// packed pair: conflict-free scalar shared loads whatever the rows' 4-byte misalignment).  Expanded form with exact
// recomputation (see distance_tlp_kernel): 80 FFMA2 per 4 channels and 64 pixels, so a (tile, all K) unit costs the
// FMA pipe ~86 us at C = 19, K = 2048 -- the same as the HBM time of its 4 MB.
constexpr int kStStages = 3;

// TP pixels per tile, KC channels per stage, NW consumer warps, PX pixels per lane (TP = 32 NW PX)
template <int CT, int TP, int KC, int NW, int PX>
struct StreamCfg {
    static_assert(TP == 32 * NW * PX && KC % 4 == 0 && KC <= 32 && PX % 2 == 0, "stream tile shape");
    static constexpr int kRS = TP + 8;                      // row stride in floats: head misalignment (3) + 16-byte round-up (3)
    static constexpr int kThreads = (NW + 1) * 32;
    static constexpr int stage_floats = KC * kRS + CT * KC;
    static constexpr int bar_off = kStStages * stage_floats * 4;
    static constexpr int pn_off = bar_off + 2 * kStStages * 8;
    static constexpr int total = pn_off + CT * 4 + 16;
    static_assert((stage_floats * 4) % 16 == 0 && total <= 227 * 1024, "stream shared memory");
};

template <int CT, int TP, int KC, int NW, int PX>
__global__ void __launch_bounds__(StreamCfg<CT, TP, KC, NW, PX>::kThreads, 1) distance_stream_kernel(const DistArgs a, int64_t units) {
    using Cfg = StreamCfg<CT, TP, KC, NW, PX>;
    constexpr int RS = Cfg::kRS, PP = PX / 2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* ring = reinterpret_cast<float*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + Cfg::bar_off);
    uint64_t* empty = full + kStStages;
    float* pn = reinterpret_cast<float*>(smem_raw + Cfg::pn_off);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t P = a.P, K = a.K;
    const int64_t T = (P + TP - 1) / TP;             // tiles per image
    const int chunks = static_cast<int>(K / KC);      // stages per unit
    if (tid == 0) {
        for (int s = 0; s < kStStages; ++s) {
            mbar_init(&full[s], 32);
            mbar_init(&empty[s], NW);
        }
        mbar_fence_init();
    }
    // the zero rows of the prototype slices (classes C..CT-1) are written once; bulk copies only touch rows < C
    for (int i = tid; i < kStStages * Cfg::stage_floats; i += Cfg::kThreads) ring[i] = 0.f;
    for (int c = warp; c < CT; c += NW + 1) {   // |p_c|^2 in fp64
        double sq = 0.0;
        if (c < a.C)
            for (int64_t k = lane; k < K; k += 32) {
                const double v = static_cast<double>(a.protos[c * K + k]);
                sq += v * v;
            }
        sq = warp_sum_d(sq);
        if (lane == 0) pn[c] = static_cast<float>(sq);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int64_t total_elems = a.N * K * P;

    if (warp == NW) {
        // ===== producer: lane l < KC stages channel row k0 + l of the tile; lanes < C also stage the prototype slice of class l
        int stage = 0;
        uint32_t phase = 0;
        for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
            const int64_t n = u / T, p0 = (u - n * T) * TP;
            const int len = static_cast<int>(min(static_cast<int64_t>(TP), P - p0));
            for (int ch = 0; ch < chunks; ++ch) {
                mbar_wait(&empty[stage], phase ^ 1u);
                float* st = ring + stage * Cfg::stage_floats;
                uint32_t nbytes = 0;
                if (lane < KC) {
                    const int64_t e0 = (n * K + static_cast<int64_t>(ch) * KC + lane) * P + p0;
                    const int o = static_cast<int>(e0 & 3);
                    const int64_t ea = e0 - o;
                    nbytes = static_cast<uint32_t>(((o + len) * 4 + 15) & ~15);
                    int patch = 0;
                    if ((ea + nbytes / 4) > total_elems) {  // never read past the end of the tensor
                        nbytes = static_cast<uint32_t>((total_elems - ea) * 4) & ~15u;
                        patch = (o + len) - static_cast<int>(nbytes / 4);
                    }
                    float* dst = st + lane * RS;
                    for (int i = 0; i < patch; ++i) dst[nbytes / 4 + i] = a.feat[ea + nbytes / 4 + i];
                    if (patch) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    const uint32_t pbytes = lane < a.C ? KC * 4 : 0;
                    mbar_arrive_expect_tx(&full[stage], nbytes + pbytes);
                    if (nbytes) bulk_g2s(dst, a.feat + ea, nbytes, &full[stage]);
                } else {
                    const uint32_t pbytes = lane < a.C ? KC * 4 : 0;
                    mbar_arrive_expect_tx(&full[stage], pbytes);
                }
                if (lane < a.C) bulk_g2s(st + KC * RS + lane * KC, a.protos + lane * K + static_cast<int64_t>(ch) * KC, KC * 4, &full[stage]);
                if (++stage == kStStages) { stage = 0; phase ^= 1u; }
            }
        }
    } else {
        // ===== consumers: warp w owns pixels [32 PX w, 32 PX (w + 1)) of the tile, lane l the pixels 32 PX w + l + 32 i
        int stage = 0;
        uint32_t phase = 0;
        const int px0 = warp * 32 * PX + lane;
        for (int64_t u = blockIdx.x; u < units; u += gridDim.x) {
            const int64_t n = u / T, p0 = (u - n * T) * TP;
            f32x2 acc[CT][PP];
#pragma unroll
            for (int c = 0; c < CT; ++c)
#pragma unroll
                for (int h = 0; h < PP; ++h) acc[c][h] = 0ull;
            f32x2 nrm[2][PP];
#pragma unroll
            for (int h = 0; h < PP; ++h) { nrm[0][h] = 0ull; nrm[1][h] = 0ull; }
            const unsigned emod = static_cast<unsigned>((n * K * P + p0) & 3), pm = static_cast<unsigned>(P & 3);
            for (int ch = 0; ch < chunks; ++ch) {
                const float* st = ring + stage * Cfg::stage_floats;
                const float* ps = st + KC * RS;
                mbar_wait(&full[stage], phase);
#pragma unroll(PX == 2 ? 2 : 1)
                for (int g = 0; g < KC / 4; ++g) {
                    f32x2 xp[4][PP];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const unsigned k = static_cast<unsigned>(ch * KC + g * 4 + q);
                        const unsigned o = (emod + k * pm) & 3u;            // this row's head misalignment
                        const float* row = st + (g * 4 + q) * RS + o + px0;
#pragma unroll
                        for (int h = 0; h < PP; ++h) xp[q][h] = pack2(row[64 * h], row[64 * h + 32]);
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q)
#pragma unroll
                        for (int h = 0; h < PP; ++h) nrm[q & 1][h] = fma2(xp[q][h], xp[q][h], nrm[q & 1][h]);
#pragma unroll
                    for (int c = 0; c < CT; ++c) {
                        const float4 pr = *reinterpret_cast<const float4*>(ps + c * KC + g * 4);
                        const float pv[4] = {pr.x, pr.y, pr.z, pr.w};
#pragma unroll
                        for (int q = 0; q < 4; ++q)
#pragma unroll
                            for (int h = 0; h < PP; ++h) acc[c][h] = fma2(pack2(pv[q], pv[q]), xp[q][h], acc[c][h]);
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[stage]);
                if (++stage == kStStages) { stage = 0; phase ^= 1u; }
            }
            // ---- unit finished: d^2 = |f|^2 + |p|^2 - 2 f.p, exact recomputation inside the cancellation region
            float d2[PX][CT];
            unsigned fl[PX];
            unsigned any = 0;
#pragma unroll
            for (int i = 0; i < PX; ++i) {
                const int64_t p = p0 + px0 + 32 * i;
                float f0, f1;
                unpack2(add2(nrm[0][i >> 1], nrm[1][i >> 1]), f0, f1);
                const float F = (i & 1) ? f1 : f0;
                fl[i] = 0;
#pragma unroll
                for (int c = 0; c < CT; ++c) {
                    float lo, hi;
                    unpack2(acc[c][i >> 1], lo, hi);
                    const float scale = F + pn[c];
                    d2[i][c] = fmaf(-2.0f, (i & 1) ? hi : lo, scale);
                    if (p < P && c < a.C && d2[i][c] < kExpTheta * scale) fl[i] |= 1u << c;
                }
                any |= fl[i];
            }
            if (__any_sync(0xffffffffu, any != 0)) {
#pragma unroll
                for (int i = 0; i < PX; ++i) {
                    unsigned pend = __ballot_sync(0xffffffffu, fl[i] != 0);
                    while (pend) {
                        const int src = __ffs(pend) - 1;
                        pend &= pend - 1;
                        const unsigned m = __shfl_sync(0xffffffffu, fl[i], src);
                        const int pp = __shfl_sync(0xffffffffu, static_cast<int>(p0 + px0 + 32 * i), src);
                        float ex[CT];
                        exact_d2_pixel<CT>(a, n, pp, m, ex);
                        if (lane == src) {
#pragma unroll
                            for (int c = 0; c < CT; ++c)
                                if ((m >> c) & 1u) d2[i][c] = ex[c];
                        }
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < PX; ++i) {
                const int64_t p = p0 + px0 + 32 * i;
                if (p < P) {
                    float d[CT];
#pragma unroll
                    for (int c = 0; c < CT; ++c) d[c] = sqrtf(fmaxf(d2[i][c], 0.f));
                    distance_epilogue<CT>(a, d, n, p);
                }
            }
        }
    }
}

template <int CT, int TP, int KC, int NW, int PX>
static int launch_distance_stream_cfg(DistArgs& a, cudaStream_t st) {
    using Cfg = StreamCfg<CT, TP, KC, NW, PX>;
    cudaError_t e = cudaFuncSetAttribute(distance_stream_kernel<CT, TP, KC, NW, PX>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::total);
    if (e != cudaSuccess) return cuda_status(e, "tss_proto_distance: smem attribute");
    const int64_t units = a.N * ((a.P + TP - 1) / TP);
    const unsigned grid = static_cast<unsigned>(units < kNumSMs ? units : kNumSMs);
    distance_stream_kernel<CT, TP, KC, NW, PX><<<grid, Cfg::kThreads, Cfg::total, st>>>(a, units);
    TSS_LAUNCH_CHECK("tss_proto_distance(stream)");
    return TSS_OK;
}

// shape: 0 = automatic, 1 = 512 px x 32 ch, 8 warps x 2 px; 2 = 512 x 32, 4 warps x 4 px; 3 = 1024 x 16, 8 warps x 4 px;
// 4 = 1024 x 16, 16 warps x 2 px
template <int CT>
static int launch_distance_stream(DistArgs& a, int shape, cudaStream_t st) {
    if (shape == 0) {
        // the 1024-pixel tiles (twice the warps or pixels per lane behind each prototype load) need enough tiles to fill
        // the machine; small batches keep the 512-pixel tile
        const int64_t big = a.N * ((a.P + 1023) / 1024);
        shape = big >= (kNumSMs * 7) / 8 ? TSS_STREAM_AUTO_BIG : 1;
    }
    switch (shape) {
#ifdef TSS_STREAM_ALL_SHAPES   // the two shapes that lost the sweep (profiles/r02_distance_forms.txt) are built on request only
        case 2: return launch_distance_stream_cfg<CT, 512, 32, 4, 4>(a, st);
        case 4: return launch_distance_stream_cfg<CT, 1024, 16, 16, 2>(a, st);
#endif
        case 3: return launch_distance_stream_cfg<CT, 1024, 16, 8, 4>(a, st);
        default: return launch_distance_stream_cfg<CT, 512, 32, 8, 2>(a, st);
    }
}

// Relative cost of one (PX, SI) form on this problem: rounds of the persistent loop on the busiest warp x channel
// steps per round (+ a per-unit constant) x pixels a round retires per SM x the measured per-pixel-channel cost of
// the form (4 px per lane holds fewer warps: x1.05; 8-lane parts fetch 32-byte pieces of a row: x1.05).  B200
// measurements behind the factors: profiles/r01_distance_forms.txt.
template <int CT, int PX, int PF>
static double tlp_cost(const DistArgs& a, int SI) {
    constexpr int W = TlpCfg<CT, PX, PF>::kWarps;
    const int unit_px = PX * 32 / SI;
    const int64_t units = a.N * ((a.P + unit_px - 1) / unit_px);
    const int64_t slots = static_cast<int64_t>(kNumSMs) * W;
    const double rounds = static_cast<double>((units + slots - 1) / slots);
    double f = PF == 0 ? 1.0 : PF == 1 ? (PX == 4 ? 0.87 : 0.93) : (PX == 4 ? 0.85 : 0.88);
    if (SI == 4) f *= 1.05;
    return rounds * (static_cast<double>(a.K) / SI + 16.0) * (W * PX * 32) * f;
}

template <int CT, int PX, int PF, bool EXP>
static int launch_distance_tlp(DistArgs& a, int SI, cudaStream_t st) {
    constexpr int W = TlpCfg<CT, PX, PF>::kWarps;
    const size_t smem = sizeof(float) * (CT * static_cast<size_t>(a.K + 4 * SI) + CT);
    const int unit_px = PX * 32 / SI;
    const int64_t units = a.N * ((a.P + unit_px - 1) / unit_px);
    cudaError_t e = cudaFuncSetAttribute(distance_tlp_kernel<CT, PX, PF, EXP>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    if (e != cudaSuccess) return cuda_status(e, "tss_proto_distance: smem attribute");
    const int64_t ctas = (units + W - 1) / W;
    const unsigned grid = static_cast<unsigned>(ctas < kNumSMs ? ctas : kNumSMs);
    distance_tlp_kernel<CT, PX, PF, EXP><<<grid, TlpCfg<CT, PX, PF>::kThreads, smem, st>>>(a, units, SI);
    TSS_LAUNCH_CHECK("tss_proto_distance(tlp)");
    return TSS_OK;
}

// `form` of tss_proto_distance (0 = everything automatic), see include/tss_b200.h
struct DistForm {
    int variant;    // 0 automatic, 1 = 8-warp register-pipeline kernel, 2 / 4 = thread-level-parallel kernel with 2 / 4 px per lane, 8 = streaming kernel
    int prefetch;   // -1 automatic, 0 / 1 / 2 = register prefetch depth in 4-channel groups
    int split;      // 0 automatic, else the K split
    int algo;       // 0 automatic, 1 direct difference, 2 expanded with exact recomputation
};
static DistForm decode_form(int form) {
    DistForm f;
    f.variant = form & 0xf;
    f.prefetch = ((form >> 4) & 0xf) - 1;
    f.split = (form >> 8) & 0xf;
    f.algo = (form >> 12) & 0xf;
    return f;
}

template <int CT, bool EXP>
static int launch_distance_tlp_form(DistArgs& a, int px, int pf, int si, cudaStream_t st) {
    if (px == 2) {
        if (pf == 0) return launch_distance_tlp<CT, 2, 0, EXP>(a, si, st);
        if (pf == 1) return launch_distance_tlp<CT, 2, 1, EXP>(a, si, st);
        return launch_distance_tlp<CT, 2, 2, EXP>(a, si, st);
    }
    return pf ? launch_distance_tlp<CT, 4, 1, EXP>(a, si, st) : launch_distance_tlp<CT, 4, 0, EXP>(a, si, st);
}

// returns TSS_OK after a launch, or -1 when the thread-level-parallel form does not apply
template <int CT>
static int launch_distance_auto(DistArgs& a, const DistForm& fm, cudaStream_t st) {
    if (fm.variant == 1) return -1;
    if (sizeof(float) * (CT * static_cast<size_t>(a.K + 16) + CT) > 226 * 1024) return -1;   // prototypes must be resident
    int best_px = 2, best_si = 1, best_pf = 0;
    double best = 1e300;
    for (int px = 2; px <= 4; px += 2) {
        if (fm.variant != 0 && fm.variant != px) continue;
        for (int pf = 0; pf <= (px == 2 ? 2 : 1); ++pf) {   // depth 2 only with 2 px per lane (registers)
            if (fm.prefetch >= 0 && fm.prefetch != pf) continue;
            for (int si = 1; si <= 4; si *= 2) {
                if (fm.split != 0 && fm.split != si) continue;
                double c;
                if (px == 2) c = pf == 0 ? tlp_cost<CT, 2, 0>(a, si) : pf == 1 ? tlp_cost<CT, 2, 1>(a, si) : tlp_cost<CT, 2, 2>(a, si);
                else c = pf == 0 ? tlp_cost<CT, 4, 0>(a, si) : tlp_cost<CT, 4, 1>(a, si);
                if (c < best) { best = c; best_px = px; best_si = si; best_pf = pf; }
            }
        }
    }
    if (best == 1e300) return -1;   // forced split 8 / unsupported combination: the register-pipeline kernel
    // expanded form: half the FP32 work; worth it once the K loop dominates the per-unit overhead
    const bool expanded = fm.algo == 2 || (fm.algo == 0 && a.K >= 128);
    return expanded ? launch_distance_tlp_form<CT, true>(a, best_px, best_pf, best_si, st)
                    : launch_distance_tlp_form<CT, false>(a, best_px, best_pf, best_si, st);
}

template <int CT>
static int launch_distance(DistArgs& a, int form, cudaStream_t st) {
    const DistForm fm = decode_form(form);
    const bool aligned = ((reinterpret_cast<uintptr_t>(a.feat) | reinterpret_cast<uintptr_t>(a.protos)) & 15) == 0;   // bulk copies
    if (aligned && (fm.variant == 8 || (fm.variant == 0 && fm.algo != 1 && fm.prefetch < 0 && fm.split == 0 && a.K >= 512)))
        return launch_distance_stream<CT>(a, fm.variant == 8 ? fm.prefetch + 1 : 0, st);   // variant 8: `prefetch` selects the tile shape
    {
        const int r = launch_distance_auto<CT>(a, fm, st);
        if (r != -1) return r;
    }
    // whole prototype matrix resident when it fits; else the largest K chunk 32*g (g | K/32, g <= 16)
    const size_t fixed = sizeof(float) * static_cast<size_t>(kDWarps) * CT * 2 * 32;
    const size_t budget = 224 * 1024;
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
    // K split: the smallest S in {1,2,4,8} that gives every warp of the machine ~6 warp tiles (load balance)
    const int64_t groups = a.N * ((a.P + kDWarpPx - 1) / kDWarpPx);
    int S = 1;
    while (S < 8 && groups * S < 6ll * kNumSMs * kDWarps) S *= 2;
    if (fm.split > 0) S = fm.split;
    const int tile_px = kDWarpPx * (kDWarps / S);
    const int64_t tiles = a.N * ((a.P + tile_px - 1) / tile_px);
    const unsigned grid = static_cast<unsigned>(tiles < kNumSMs ? tiles : kNumSMs);
    distance_kernel<CT><<<grid, kDThreads, smem, st>>>(a, tiles, S);
    TSS_LAUNCH_CHECK("tss_proto_distance");
    return TSS_OK;
}

// ------------------------------------------------------------------------------------------------
// Distance / weights / rectified labels straight from the network's 1/8-stride features: the
// reference first interpolates feat to a quarter of the full image (self_training.py:69), then full2weak resizes
// that map per image, crops and mirrors it (:145-158), and only then takes distances -- two materialised
// K-channel tensors.  Here a thread owns one output pixel and evaluates BOTH bilinear stages on the fly, per
// channel, with the same fp32 operation sequence as the two ATen calls (stage 1: four quarter-map values from four
// source taps each; stage 2: the resized value from those four), then accumulates the direct-difference distances
// to the C prototypes in shared memory.  HBM traffic: the 1/8-stride features once (4 K bytes per SOURCE pixel; the
// 16 taps of a pixel hit L1/L2) + soft / hard / label per output pixel.
struct SampleArgs {
    const int32_t* params;   // [N,6] {rh, rw, y0, x0, flip, 0}
    int h8, w8, Hq, Wq, TH, TW;
};

__device__ __forceinline__ void axis_tap(float scale, int dst, int n_in, int& i0, int& i1, float& l0, float& l1) {
    const float s = __fmul_rn(scale, static_cast<float>(dst));   // ATen: src = scale * dst, all fp32
    i0 = min(static_cast<int>(s), n_in - 1);
    i1 = i0 + (i0 < n_in - 1 ? 1 : 0);
    l1 = __fsub_rn(s, static_cast<float>(i0));
    l0 = __fsub_rn(1.0f, l1);
}
__device__ __forceinline__ float axis_scale(int n_in, int n_out) {
    return n_out > 1 ? static_cast<float>(n_in - 1) / static_cast<float>(n_out - 1) : 0.f;
}

constexpr int kSampThreads = 128;

// KS = channels sampled per step (their 16 KS tap loads are in flight together): 4 at 241 registers (2 CTAs per SM), 2
// under a 128-register cap (4 CTAs per SM)
template <int CT, int KS>
__global__ void __launch_bounds__(kSampThreads, KS == 4 ? 1 : 4) distance_sampled_kernel(const DistArgs a, const SampleArgs sp) {
    extern __shared__ __align__(16) float ps[];   // [CT][K + 4]
    const int K = static_cast<int>(a.K), PS = K + 4;
    for (int i = threadIdx.x; i < CT * K; i += kSampThreads) {
        const int c = i / K, k = i - c * K;
        ps[c * PS + k] = (c < a.C) ? a.protos[i] : 0.f;
    }
    __syncthreads();
    const int64_t n = blockIdx.y;
    const int p = blockIdx.x * kSampThreads + threadIdx.x;
    if (p >= sp.TH * sp.TW) return;
    const int y = p / sp.TW, x = p - y * sp.TW;
    const int rh = sp.params[n * 6 + 0], rw = sp.params[n * 6 + 1], y0 = sp.params[n * 6 + 2], x0 = sp.params[n * 6 + 3];
    const int flip = sp.params[n * 6 + 4];
    // stage 2 (full2weak): taps of the resized-and-cropped pixel in the quarter map
    int Y[2], X[2];
    float ly0, ly1, lx0, lx1;
    axis_tap(axis_scale(sp.Hq, rh), y0 + y, sp.Hq, Y[0], Y[1], ly0, ly1);
    axis_tap(axis_scale(sp.Wq, rw), x0 + (flip ? sp.TW - 1 - x : x), sp.Wq, X[0], X[1], lx0, lx1);
    // stage 1 (1/8 -> 1/4 interpolation): taps of those quarter-map points in the source map
    const float s1y = axis_scale(sp.h8, sp.Hq), s1x = axis_scale(sp.w8, sp.Wq);
    int off[2][2][2][2];   // [quarter row a][source row r][quarter col b][source col s]
    float my0[2], my1[2], mx0[2], mx1[2];
    {
        int iy[2][2], jx[2][2];
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            axis_tap(s1y, Y[t], sp.h8, iy[t][0], iy[t][1], my0[t], my1[t]);
            axis_tap(s1x, X[t], sp.w8, jx[t][0], jx[t][1], mx0[t], mx1[t]);
        }
#pragma unroll
        for (int t = 0; t < 2; ++t)
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int b = 0; b < 2; ++b)
#pragma unroll
                    for (int q = 0; q < 2; ++q) off[t][r][b][q] = iy[t][r] * sp.w8 + jx[b][q];
    }
    const int64_t plane8 = static_cast<int64_t>(sp.h8) * sp.w8;
    const float* f = a.feat + n * a.K * plane8;
    float acc[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) acc[c] = 0.f;
    for (int k0 = 0; k0 < K; k0 += 4) {
        float v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (KS == 2 && q == 2) asm volatile("" ::: "memory");   // keep the two halves' loads apart (register cap)
            const float* ch = f + (k0 + q) * plane8;
            float Q[2][2];
#pragma unroll
            for (int t = 0; t < 2; ++t)
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const float top = fmaf(mx1[b], __ldg(ch + off[t][0][b][1]), __fmul_rn(mx0[b], __ldg(ch + off[t][0][b][0])));
                    const float bot = fmaf(mx1[b], __ldg(ch + off[t][1][b][1]), __fmul_rn(mx0[b], __ldg(ch + off[t][1][b][0])));
                    Q[t][b] = fmaf(my1[t], bot, __fmul_rn(my0[t], top));
                }
            const float top = fmaf(lx1, Q[0][1], __fmul_rn(lx0, Q[0][0]));
            const float bot = fmaf(lx1, Q[1][1], __fmul_rn(lx0, Q[1][0]));
            v[q] = fmaf(ly1, bot, __fmul_rn(ly0, top));
        }
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            const float4 pr = *reinterpret_cast<const float4*>(ps + c * PS + k0);
            float d;
            d = __fsub_rn(pr.x, v[0]); acc[c] = fmaf(d, d, acc[c]);
            d = __fsub_rn(pr.y, v[1]); acc[c] = fmaf(d, d, acc[c]);
            d = __fsub_rn(pr.z, v[2]); acc[c] = fmaf(d, d, acc[c]);
            d = __fsub_rn(pr.w, v[3]); acc[c] = fmaf(d, d, acc[c]);
        }
    }
    float d[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) d[c] = sqrtf(acc[c]);
    distance_epilogue<CT>(a, d, n, p);
}

template <int CT>
static int launch_distance_sampled(const DistArgs& a, const SampleArgs& sp, int flags, cudaStream_t st) {
    const size_t smem = sizeof(float) * CT * static_cast<size_t>(a.K + 4);
    TSS_REQUIRE(smem <= 226 * 1024, TSS_E_ARG, "tss_proto_distance_sampled: %d x %lld prototypes do not fit shared memory", CT,
                static_cast<long long>(a.K));
    dim3 grid(static_cast<unsigned>((sp.TH * sp.TW + kSampThreads - 1) / kSampThreads), static_cast<unsigned>(a.N));
    cudaError_t e;
    if (flags & 1) {
        e = cudaFuncSetAttribute(distance_sampled_kernel<CT, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_status(e, "tss_proto_distance_sampled: smem attribute");
        distance_sampled_kernel<CT, 4><<<grid, kSampThreads, smem, st>>>(a, sp);
    } else {
        e = cudaFuncSetAttribute(distance_sampled_kernel<CT, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_status(e, "tss_proto_distance_sampled: smem attribute");
        distance_sampled_kernel<CT, 2><<<grid, kSampThreads, smem, st>>>(a, sp);
    }
    TSS_LAUNCH_CHECK("tss_proto_distance_sampled");
    return TSS_OK;
}

}  // namespace tss

extern "C" int tss_proto_distance_sampled(const float* feat8, const float* protos, int64_t n_img, int64_t channels, int h8,
                                          int w8, int Hq, int Wq, const int32_t* params, int TH, int TW, int num_classes,
                                          float temperature, float* dist_out, float* weight_out, const float* soft,
                                          const void* hard, int hard_dtype, float train_thred, int ignore_index,
                                          int64_t* label_out, int flags, void* stream) {
    using namespace tss;
    TSS_REQUIRE(feat8 && protos && params, TSS_E_ARG, "tss_proto_distance_sampled: null pointer");
    TSS_REQUIRE(flags == 0 || flags == 1, TSS_E_ARG, "tss_proto_distance_sampled: bad flags %d", flags);
    TSS_REQUIRE(n_img >= 0 && n_img <= 65535 && channels > 0 && h8 > 0 && w8 > 0 && Hq > 0 && Wq > 0 && TH > 0 && TW > 0, TSS_E_ARG,
                "tss_proto_distance_sampled: bad extents");
    TSS_REQUIRE(channels % 4 == 0, TSS_E_ARG, "tss_proto_distance_sampled: channels (%lld) must be a multiple of 4", static_cast<long long>(channels));
    TSS_REQUIRE(num_classes >= 1 && num_classes <= 20, TSS_E_CLASSES, "tss_proto_distance_sampled: num_classes %d outside [1, 20]", num_classes);
    TSS_REQUIRE(dist_out || weight_out || (soft && label_out), TSS_E_ARG, "tss_proto_distance_sampled: no output requested");
    TSS_REQUIRE(hard == nullptr || hard_dtype == TSS_U8 || hard_dtype == TSS_I32 || hard_dtype == TSS_I64, TSS_E_ARG, "tss_proto_distance_sampled: bad hard dtype");
    if (n_img == 0) return TSS_OK;
    DistArgs a;
    a.feat = feat8; a.protos = protos; a.dist_out = dist_out; a.weight_out = weight_out; a.soft = soft;
    a.hard = hard; a.label_out = label_out; a.N = n_img; a.K = channels; a.P = static_cast<int64_t>(TH) * TW; a.C = num_classes;
    a.temperature = temperature; a.train_thred = train_thred; a.hard_dtype = hard_dtype; a.ignore = ignore_index;
    a.KC = 0;
    SampleArgs sp{params, h8, w8, Hq, Wq, TH, TW};
    cudaStream_t st = as_stream(stream);
    // class-count specialisations: Freiburg (13), Cityscapes (19), and 20; smaller counts run the next one up (guarded)
    if (num_classes <= 13) return launch_distance_sampled<13>(a, sp, flags, st);
    if (num_classes <= 19) return launch_distance_sampled<19>(a, sp, flags, st);
    return launch_distance_sampled<20>(a, sp, flags, st);
}

extern "C" int tss_proto_distance(const float* feat, const float* protos, int64_t n_img, int64_t channels,
                                  int64_t plane, int num_classes, float temperature, float* dist_out,
                                  float* weight_out, const float* soft, const void* hard, int hard_dtype,
                                  float train_thred, int ignore_index, int64_t* label_out, int form, void* stream) {
    using namespace tss;
    TSS_REQUIRE(feat && protos, TSS_E_ARG, "tss_proto_distance: null pointer");
    TSS_REQUIRE(n_img >= 0 && plane > 0 && plane < (1ll << 31) && channels > 0, TSS_E_ARG, "tss_proto_distance: bad extents");
    {
        const DistForm fm = decode_form(form);
        TSS_REQUIRE(form >= 0 && form < (1 << 16) && (fm.variant == 0 || fm.variant == 1 || fm.variant == 2 || fm.variant == 4 || fm.variant == 8) &&
                    fm.prefetch <= (fm.variant == 8 ? 3 : 2) && (fm.split == 0 || fm.split == 1 || fm.split == 2 || fm.split == 4 || fm.split == 8) &&
                    fm.algo <= 2, TSS_E_ARG, "tss_proto_distance: bad form 0x%x", form);
    }
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
    // class-count specialisations: Freiburg (13), Cityscapes (19), and 20; smaller counts run the next one up (guarded)
    if (num_classes <= 13) return launch_distance<13>(a, form, st);
    if (num_classes <= 19) return launch_distance<19>(a, form, st);
    return launch_distance<20>(a, form, st);
}
