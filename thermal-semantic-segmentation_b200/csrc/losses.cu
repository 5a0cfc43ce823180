// A11-A13: cross-entropy with ignore_index, reverse cross-entropy, focal loss -- forward sums and
// input gradients (self_training.py:27,114-118,167-174; utils/loss.py:142-184).
// HBM-bound: each logit is read once (4C + label bytes per pixel), gradients written once.
// Per-pixel terms are accumulated in fp64 per thread -> per CTA -> a fixed-order final reduction, so the
// loss is bit-reproducible run to run.
#include "common.cuh"

namespace tss {

constexpr int kLossThreads = 256;
constexpr int kLossMaxBlocks = kNumSMs * 8;
#ifndef TSS_LOSS_FWD_CTAS_PER_SM
#define TSS_LOSS_FWD_CTAS_PER_SM 2
#endif

template <int CT, bool EXACT>
struct PixelSoftmax {
    float p[CT];   // exp(x - max), later normalised on demand
    float m, s;
    float xy;      // raw logit of class y (picked from registers: the streaming loads bypass L1)
    __device__ __forceinline__ void load(const float* x, int Crt, int64_t plane, int y = -1) {
        const int C = EXACT ? CT : Crt;   // EXACT: the class count is the template constant, every guard folds
        // unconditional loads (class index clamped when not EXACT): all C loads of the pixel are in flight together
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            p[c] = ldg_stream_f(x);
            if (EXACT || c + 1 < C) x += plane;
        }
        finish(Crt, y);
    }
    // p[] holds the raw logits (registers): max, the label's logit, exponentials and their sum
    __device__ __forceinline__ void finish(int Crt, int y = -1) {
        const int C = EXACT ? CT : Crt;
        m = -INFINITY;
        xy = 0.f;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < C) {
                m = fmaxf(m, p[c]);
                xy = (c == y) ? p[c] : xy;
            }
        // exp(x - m) = 2^((x - m) * log2 e): one FFMA + one MUFU.EX2 per class (2 ulp; the loss bar is 1e-5 relative)
        const float mb = m * kLog2e;
        s = 0.f;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < C) {
                p[c] = ex2_approx(fmaf(p[c], kLog2e, -mb));
                s += p[c];
            }
    }
    __device__ __forceinline__ float log_s() const { return __logf(s); }
    __device__ __forceinline__ float inv_s() const { return __frcp_rn(s); }
};

template <typename T>
__device__ __forceinline__ long long target_at(const void* t, int64_t i) { return load_label<T>(t, i); }

// per-CTA partials: kAcc doubles per block (0 CE sum, 1 RCE sum, 2 pixels with a label in [0, C), 3 labels the
// reference raises on, 4 pixels labelled exactly C -- one_hot(label, C+1)[:, :C] accepts those in SelfTrain.rce)
constexpr int kAcc = 5;
constexpr int kAccStride = 8;

__device__ __forceinline__ void block_reduce(double (&v)[kAcc], double* out) {
    __shared__ double sh[kAcc][kLossThreads / 32];
#pragma unroll
    for (int i = 0; i < kAcc; ++i) v[i] = warp_sum_d(v[i]);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
#pragma unroll
        for (int i = 0; i < kAcc; ++i) sh[i][warp] = v[i];
    __syncthreads();
    if (threadIdx.x < kAcc) {
        double t = 0.0;
        for (int w = 0; w < kLossThreads / 32; ++w) t += sh[threadIdx.x][w];
        out[threadIdx.x] = t;
    }
}

// losses (or NULL): {alpha ce + beta rce, ce, rce} in fp32 with the host mirror's own arithmetic -- sums divided in
// fp64, rounded to fp32, mixed in fp32 -- so that the op needs no scalar follow-up kernels
struct LossMix {
    float* losses;
    float alpha, beta;
    int with_rce, ce_sum;
};

// Final reduction of the per-CTA partials in a FIXED order (bit-reproducible), executed by whichever CTA finishes last:
// every CTA stores its partials, fences, and takes a ticket; the CTA that draws the last ticket reduces all partials
// (read past L1) and resets the ticket to zero for the next call.  One launch instead of two.
__device__ __forceinline__ void finish_if_last(const double* __restrict__ part, unsigned int* ticket, double* __restrict__ out,
                                               const LossMix& mix) {
    __shared__ bool last;
    __shared__ double sh[kAcc][kLossThreads / 32];
    __shared__ double tot[kAcc];
    __threadfence();     // this CTA's partials are visible device-wide ...
    __syncthreads();     // ... for every thread that stored one, before the ticket is drawn
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    const int blocks = static_cast<int>(gridDim.x);
    double v[kAcc] = {0, 0, 0, 0, 0};
    for (int b = threadIdx.x; b < blocks; b += kLossThreads)
#pragma unroll
        for (int i = 0; i < kAcc; ++i) v[i] += __ldcg(part + b * kAccStride + i);
#pragma unroll
    for (int i = 0; i < kAcc; ++i) v[i] = warp_sum_d(v[i]);
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int i = 0; i < kAcc; ++i) sh[i][threadIdx.x >> 5] = v[i];
    __syncthreads();
    if (threadIdx.x < kAcc) {
        double t = 0.0;
        for (int w = 0; w < kLossThreads / 32; ++w) t += sh[threadIdx.x][w];
        out[threadIdx.x] = t;
        tot[threadIdx.x] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *ticket = 0u;
        if (mix.losses != nullptr) {
            const double valid = tot[2];
            const float ce = static_cast<float>(mix.ce_sum ? tot[0] : tot[0] / valid);   // mean: nan when nothing is valid, as torch
            const float rce = static_cast<float>(tot[1] / (valid + tot[4] + 1e-6));      // the rce mask is label != ignore
            float t = __fmul_rn(mix.alpha, ce);
            if (mix.with_rce) t = __fadd_rn(t, __fmul_rn(mix.beta, rce));
            mix.losses[0] = t;
            mix.losses[1] = ce;
            mix.losses[2] = rce;
        }
    }
}

// -log(clamp(0, 1e-4, 1)) in fp32, as torch.log(torch.clamp(one_hot, min=1e-4)) yields off the label
__device__ __forceinline__ float rce_log_floor() { return -logf(1e-4f); }

// one pixel's CE / RCE terms from a finished PixelSoftmax.  A label equal to C is out of bounds for CrossEntropyLoss
// but inside one_hot(label, C+1): its clamped one-hot row is all 1e-4, so the pixel adds A * sum_c clamp(p_c) to the
// RCE sum and counts in the RCE mask (self_training.py:169-173); it is tallied separately in acc[4].
template <int CT, bool EXACT>
__device__ __forceinline__ void ce_rce_pixel(const PixelSoftmax<CT, EXACT>& sm, long long y, int C, int ignore, int with_rce,
                                             float A, double (&acc)[kAcc]) {
    if (y == ignore) return;
    if (y < 0 || y > C || (y == C && !with_rce)) { acc[3] += 1.0; return; }
    const int Cc = EXACT ? CT : C;
    if (y < C) {
        // -log_softmax[y] = log(s) - (x_y - m)
        acc[0] += static_cast<double>(sm.log_s() - (sm.xy - sm.m));
        acc[2] += 1.0;
    } else {
        acc[4] += 1.0;
    }
    if (with_rce) {
        const float inv = sm.inv_s();
        float r = 0.f;
        int yi = static_cast<int>(y);
        asm("" : "+r"(yi));   // keep the C class tests 32-bit (the compiler otherwise folds them back into 64-bit compares of y)
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < Cc && c != yi) r += fminf(fmaxf(sm.p[c] * inv, 1e-7f), 1.0f);
        acc[1] += static_cast<double>(A * r);
    }
}

// VEC: four consecutive pixels per thread and iteration (plane % 4 == 0, 16-byte aligned logits): C 128-bit loads
// in flight per thread instead of C 32-bit ones -- the scalar form cannot keep enough bytes in flight for HBM.
// Label and logits are loaded together (the logit loads do not wait for the ignore test; ignored pixels cost the
// same bytes either way, their sectors hold valid neighbours).
template <int CT, bool EXACT, typename T, bool VEC>
__global__ void __launch_bounds__(kLossThreads)
ce_rce_fwd_kernel(const float* __restrict__ logits, const void* __restrict__ target, int64_t N, int C, int64_t P,
                  int ignore, int with_rce, double* __restrict__ part, unsigned int* ticket, double* __restrict__ out, const LossMix mix) {
    double acc[kAcc] = {0, 0, 0, 0, 0};
    const float A = rce_log_floor();
    const int64_t first = blockIdx.x * static_cast<int64_t>(kLossThreads) + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kLossThreads;
    if (VEC) {
        // Four pixels per thread as two packed pairs (FFMA2 / FADD2 halve the issue slots of the exponent arguments and
        // the sums).  The label's own logit is RE-LOADED from the line that has just streamed through L2 (one 4-byte
        // load per pixel) instead of being picked out of the C registers with C compare + select pairs, and the RCE
        // sum over c != y is the sum over all classes minus the label's term (the same ex2 of the same argument, so the
        // difference carries only the rounding of the sum).  The per-pixel terms of the four pixels are added in fp32
        // and enter the fp64 accumulators once per iteration.
        const int Cc = EXACT ? CT : C;
        const int64_t Pq = P >> 2, total = N * Pq;
        const f32x2 l2e = pack2(kLog2e, kLog2e);
        PixelWalk w(first, stride, Pq);
        for (int64_t q = first; q < total; q += stride, w.next()) {
            const float* x0 = logits + w.n * C * P + 4 * w.p;
            const float* x = x0;
            float4 v[CT];
#pragma unroll
            for (int c = 0; c < CT; ++c) {
                v[c] = ldg_stream_f4(x);
                if (EXACT || c + 1 < C) x += P;
            }
            long long y[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) y[j] = target_at<T>(target, 4 * q + j);
            // label classification (32-bit tests) and the reload of the label's logit
            float xy[4];
            bool in_range[4], over[4], bad[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool ign = y[j] == ignore;
                in_range[j] = !ign && static_cast<unsigned long long>(y[j]) < static_cast<unsigned long long>(C);
                over[j] = !ign && y[j] == C && with_rce;
                bad[j] = !ign && !in_range[j] && !over[j];
                const int yc = in_range[j] ? static_cast<int>(y[j]) : 0;
                xy[j] = __ldg(x0 + static_cast<int64_t>(yc) * P + j);
            }
            float m[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) m[j] = -INFINITY;
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (c < Cc) {
                    m[0] = fmaxf(m[0], v[c].x); m[1] = fmaxf(m[1], v[c].y);
                    m[2] = fmaxf(m[2], v[c].z); m[3] = fmaxf(m[3], v[c].w);
                }
            const f32x2 nm01 = pack2(-(m[0] * kLog2e), -(m[1] * kLog2e)), nm23 = pack2(-(m[2] * kLog2e), -(m[3] * kLog2e));
            f32x2 s01 = 0ull, s23 = 0ull;
            float e[CT][4];
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (c < Cc) {
                    float a0, a1, a2, a3;
                    unpack2(fma2(pack2(v[c].x, v[c].y), l2e, nm01), a0, a1);
                    unpack2(fma2(pack2(v[c].z, v[c].w), l2e, nm23), a2, a3);
                    e[c][0] = ex2_approx(a0); e[c][1] = ex2_approx(a1); e[c][2] = ex2_approx(a2); e[c][3] = ex2_approx(a3);
                    s01 = add2(s01, pack2(e[c][0], e[c][1]));
                    s23 = add2(s23, pack2(e[c][2], e[c][3]));
                }
            float sv[4];
            unpack2(s01, sv[0], sv[1]);
            unpack2(s23, sv[2], sv[3]);
            float rsum[4] = {0.f, 0.f, 0.f, 0.f};
            if (with_rce) {
                // sum_c clamp(p_c, 1e-7, 1) = inv * sum_c max(e_c, 1e-7 s)   (p_c <= 1 up to rounding: the upper clamp is
                // a 1-ulp effect on a term of order one)
                float t[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) t[j] = 1e-7f * sv[j];
                f32x2 r01 = 0ull, r23 = 0ull;
#pragma unroll
                for (int c = 0; c < CT; ++c)
                    if (c < Cc) {
                        r01 = add2(r01, pack2(fmaxf(e[c][0], t[0]), fmaxf(e[c][1], t[1])));
                        r23 = add2(r23, pack2(fmaxf(e[c][2], t[2]), fmaxf(e[c][3], t[3])));
                    }
                unpack2(r01, rsum[0], rsum[1]);
                unpack2(r23, rsum[2], rsum[3]);
            }
            float ce4 = 0.f, rce4 = 0.f;
            int nvalid = 0, nover = 0, nbad = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                nbad += bad[j] ? 1 : 0;
                if (in_range[j]) {
                    ce4 += __logf(sv[j]) - (xy[j] - m[j]);                       // -log_softmax[y]
                    ++nvalid;
                }
                if (with_rce && (in_range[j] || over[j])) {
                    const float inv = __frcp_rn(sv[j]);
                    float own = 0.f;                                             // the label's own clamped term
                    if (in_range[j]) own = fmaxf(ex2_approx(fmaf(xy[j], kLog2e, -(m[j] * kLog2e))), 1e-7f * sv[j]);
                    rce4 += A * ((rsum[j] - own) * inv);
                    nover += over[j] ? 1 : 0;
                }
            }
            acc[0] += static_cast<double>(ce4);
            acc[1] += static_cast<double>(rce4);
            acc[2] += static_cast<double>(nvalid);
            acc[3] += static_cast<double>(nbad);
            acc[4] += static_cast<double>(nover);
        }
    } else {
        const int64_t total = N * P;
        PixelWalk w(first, stride, P);
        for (int64_t i = first; i < total; i += stride, w.next()) {
            const long long y = target_at<T>(target, i);
            PixelSoftmax<CT, EXACT> sm;
            sm.load(logits + w.n * C * P + w.p, C, P, static_cast<int>(y));
            ce_rce_pixel<CT, EXACT>(sm, y, C, ignore, with_rce, A, acc);
        }
    }
    block_reduce(acc, part + blockIdx.x * kAccStride);
    finish_if_last(part, ticket, out, mix);
}

// upstream gradients {g_total, g_ce, g_rce} (DEVICE) + the forward's sums: the backward kernel forms its own scales
struct GradMix {
    const float* grads;
    const double* fwd_out;
    float alpha, beta;
    int with_rce, ce_sum;
};

template <int CT, bool EXACT, typename T>
__global__ void __launch_bounds__(kLossThreads)
ce_rce_bwd_kernel(const float* __restrict__ logits, const void* __restrict__ target, int64_t N, int C, int64_t P,
                  int ignore, const float* __restrict__ scales, float* __restrict__ grad, const GradMix gm) {
    float ce_scale, rce_scale;
    if (gm.grads != nullptr) {
        // scales from the upstream gradients of (total, ce, rce) and the forward sums, with the host mirror's arithmetic:
        // gce = g_total alpha + g_ce, grce = g_total beta + g_rce (fp32); ce_scale = gce / valid (fp64), rce likewise
        const float gt = gm.grads[0], gce = __fadd_rn(__fmul_rn(gt, gm.alpha), gm.grads[1]);
        const float grce = __fadd_rn(gm.with_rce ? __fmul_rn(gt, gm.beta) : 0.f, gm.grads[2]);
        const double valid = gm.fwd_out[2];
        ce_scale = static_cast<float>(gm.ce_sum ? static_cast<double>(gce) : static_cast<double>(gce) / valid);
        rce_scale = static_cast<float>(static_cast<double>(grce) / (valid + gm.fwd_out[4] + 1e-6));
    } else {
        ce_scale = scales[0];
        rce_scale = scales[1];
    }
    const float A = rce_log_floor();
    const int64_t total = N * P;
    const int Cc = EXACT ? CT : C;   // per-class guards fold at compile time when EXACT
    const int64_t first = blockIdx.x * static_cast<int64_t>(kLossThreads) + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kLossThreads;
    PixelWalk w(first, stride, P);
    for (int64_t i = first; i < total; i += stride, w.next()) {
        const long long y = target_at<T>(target, i);
        float* g = grad + w.n * C * P + w.p;
        PixelSoftmax<CT, EXACT> sm;
        sm.load(logits + w.n * C * P + w.p, C, P);   // together with the label (see the forward kernel)
        if (y == ignore || y < 0 || y > C || (y == C && rce_scale == 0.f)) {
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (c < Cc) g[c * P] = 0.f;
            continue;
        }
        const float ces = y < C ? ce_scale : 0.f;   // a label equal to C carries only the RCE term (see ce_rce_pixel)
        const float inv = sm.inv_s();
        float S = 0.f;  // sum over j != y of [1e-7 <= p_j <= 1] p_j  (clamp passes gradient inside its range)
        int yi = static_cast<int>(y);
        asm("" : "+r"(yi));   // 32-bit class tests (see the forward kernel)
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < Cc) {
                sm.p[c] = sm.p[c] * inv;
                if (c != yi && sm.p[c] >= 1e-7f && sm.p[c] <= 1.0f) S += sm.p[c];
            }
        const float rs = rce_scale * A;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < Cc) {
                const float pc = sm.p[c];
                const bool is_y = c == yi;
                const float gate = (!is_y && pc >= 1e-7f && pc <= 1.0f) ? 1.f : 0.f;
                g[c * P] = ces * (pc - (is_y ? 1.f : 0.f)) + rs * pc * (gate - S);
            }
    }
}

__device__ __forceinline__ float focal_weight(float pt, float gamma) {
    const float q = 1.0f - pt;
    if (gamma == 0.f) return 1.0f;
    if (gamma == 1.f) return q;
    if (gamma == 2.f) return q * q;
    return powf(q, gamma);
}

template <int CT, bool EXACT, typename T>
__global__ void __launch_bounds__(kLossThreads)
focal_fwd_kernel(const float* __restrict__ logits, const void* __restrict__ target, int64_t N, int C, int64_t P,
                 float gamma, const float* __restrict__ alpha, double* __restrict__ part, unsigned int* ticket, double* __restrict__ out) {
    double acc[kAcc] = {0, 0, 0, 0, 0};
    const int64_t total = N * P;
    const int Cc = EXACT ? CT : C;   // per-class guards fold at compile time when EXACT
    const int64_t first = blockIdx.x * static_cast<int64_t>(kLossThreads) + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kLossThreads;
    PixelWalk w(first, stride, P);
    for (int64_t i = first; i < total; i += stride, w.next()) {
        const long long y = target_at<T>(target, i);
        PixelSoftmax<CT, EXACT> sm;
        sm.load(logits + w.n * C * P + w.p, C, P, static_cast<int>(y));
        if (y < 0 || y >= C) { acc[3] += 1.0; continue; }
        float logpt = (sm.xy - sm.m) - sm.log_s();
        const float pt = __expf(logpt);
        if (alpha) logpt *= alpha[y];
        acc[0] += static_cast<double>(-1.0f * focal_weight(pt, gamma) * logpt);
        acc[1] += 1.0;
    }
    block_reduce(acc, part + blockIdx.x * kAccStride);
    finish_if_last(part, ticket, out, LossMix{nullptr, 0.f, 0.f, 0, 0});
}

template <int CT, bool EXACT, typename T>
__global__ void __launch_bounds__(kLossThreads)
focal_bwd_kernel(const float* __restrict__ logits, const void* __restrict__ target, int64_t N, int C, int64_t P,
                 float gamma, const float* __restrict__ alpha, const float* __restrict__ scale, float* __restrict__ grad) {
    const float sc = scale[0];
    const int64_t total = N * P;
    const int Cc = EXACT ? CT : C;   // per-class guards fold at compile time when EXACT
    const int64_t first = blockIdx.x * static_cast<int64_t>(kLossThreads) + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kLossThreads;
    PixelWalk w(first, stride, P);
    for (int64_t i = first; i < total; i += stride, w.next()) {
        const long long y = target_at<T>(target, i);
        float* g = grad + w.n * C * P + w.p;
        PixelSoftmax<CT, EXACT> sm;
        sm.load(logits + w.n * C * P + w.p, C, P, static_cast<int>(y));
        if (y < 0 || y >= C) {
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (c < Cc) g[c * P] = 0.f;
            continue;
        }
        const float pt = __expf((sm.xy - sm.m) - sm.log_s());
        // pt is detached in the reference (logpt.data.exp()): d loss / d x_c = -w * alpha_y * (delta_cy - p_c)
        const float coef = -sc * focal_weight(pt, gamma) * (alpha ? alpha[y] : 1.0f);
        const float inv = sm.inv_s();
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < Cc) g[c * P] = coef * ((c == static_cast<int>(y) ? 1.f : 0.f) - sm.p[c] * inv);
    }
}

static int loss_grid(int64_t total) {
    int64_t b = (total + kLossThreads - 1) / kLossThreads;
    if (b > kLossMaxBlocks) b = kLossMaxBlocks;
    return static_cast<int>(b < 1 ? 1 : b);
}

#define TSS_DISPATCH_T(KERNEL, CT, EX, DT, GRID, ST, ...)                                                  \
    do {                                                                                                      \
        if ((DT) == TSS_U8) KERNEL<CT, EX, uint8_t><<<GRID, kLossThreads, 0, ST>>>(__VA_ARGS__);              \
        else if ((DT) == TSS_I32) KERNEL<CT, EX, int32_t><<<GRID, kLossThreads, 0, ST>>>(__VA_ARGS__);        \
        else KERNEL<CT, EX, int64_t><<<GRID, kLossThreads, 0, ST>>>(__VA_ARGS__);                             \
    } while (0)
// 13 (Freiburg) and 19 (Cityscapes) classes get exact specialisations; other counts the guarded generic form
#define TSS_DISPATCH_CT_T(KERNEL, C, DT, GRID, ST, ...)                                                       \
    do {                                                                                                      \
        if ((C) == 13) TSS_DISPATCH_T(KERNEL, 13, true, DT, GRID, ST, __VA_ARGS__);                           \
        else if ((C) == 19) TSS_DISPATCH_T(KERNEL, 19, true, DT, GRID, ST, __VA_ARGS__);                      \
        else if ((C) < 13) TSS_DISPATCH_T(KERNEL, 13, false, DT, GRID, ST, __VA_ARGS__);                      \
        else if ((C) < 19) TSS_DISPATCH_T(KERNEL, 19, false, DT, GRID, ST, __VA_ARGS__);                      \
        else TSS_DISPATCH_T(KERNEL, TSS_MAX_CLASSES, false, DT, GRID, ST, __VA_ARGS__);                       \
    } while (0)

// the same with a trailing bool template argument (VEC)
#define TSS_DISPATCH_TV(KERNEL, CT, EX, V, DT, GRID, ST, ...)                                              \
    do {                                                                                                      \
        if ((DT) == TSS_U8) KERNEL<CT, EX, uint8_t, V><<<GRID, kLossThreads, 0, ST>>>(__VA_ARGS__);           \
        else if ((DT) == TSS_I32) KERNEL<CT, EX, int32_t, V><<<GRID, kLossThreads, 0, ST>>>(__VA_ARGS__);     \
        else KERNEL<CT, EX, int64_t, V><<<GRID, kLossThreads, 0, ST>>>(__VA_ARGS__);                          \
    } while (0)
#define TSS_DISPATCH_CT_TV(KERNEL, V, C, DT, GRID, ST, ...)                                                   \
    do {                                                                                                      \
        if ((C) == 13) TSS_DISPATCH_TV(KERNEL, 13, true, V, DT, GRID, ST, __VA_ARGS__);                       \
        else if ((C) == 19) TSS_DISPATCH_TV(KERNEL, 19, true, V, DT, GRID, ST, __VA_ARGS__);                  \
        else if ((C) < 13) TSS_DISPATCH_TV(KERNEL, 13, false, V, DT, GRID, ST, __VA_ARGS__);                  \
        else if ((C) < 19) TSS_DISPATCH_TV(KERNEL, 19, false, V, DT, GRID, ST, __VA_ARGS__);                  \
        else TSS_DISPATCH_TV(KERNEL, TSS_MAX_CLASSES, false, V, DT, GRID, ST, __VA_ARGS__);                   \
    } while (0)

static int check_loss_args(const char* who, const void* logits, const void* target, int dt, int64_t N, int C, int64_t P) {
    TSS_REQUIRE(logits && target, TSS_E_ARG, "%s: null pointer", who);
    TSS_REQUIRE(N >= 0 && P > 0, TSS_E_ARG, "%s: bad extents", who);
    TSS_REQUIRE(C >= 1 && C <= TSS_MAX_CLASSES, TSS_E_CLASSES, "%s: num_classes %d", who, C);
    TSS_REQUIRE(dt == TSS_U8 || dt == TSS_I32 || dt == TSS_I64, TSS_E_ARG, "%s: bad target dtype %d", who, dt);
    return TSS_OK;
}

}  // namespace tss

extern "C" size_t tss_loss_workspace_bytes(void) { return static_cast<size_t>(tss::kLossMaxBlocks) * tss::kAccStride * sizeof(double) + 16; }
namespace tss {
static inline unsigned int* ticket_of(void* workspace) {   // the trailing 16 bytes of the workspace
    return reinterpret_cast<unsigned int*>(static_cast<char*>(workspace) + static_cast<size_t>(kLossMaxBlocks) * kAccStride * sizeof(double));
}
}  // namespace tss

static int ce_rce_fwd_impl(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                           int64_t plane, int ignore_index, int with_rce, double* out, void* workspace, void* stream,
                           const tss::LossMix mix);

extern "C" int tss_ce_rce_fwd(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                              int64_t plane, int ignore_index, int with_rce, double* out, void* workspace, void* stream) {
    return ce_rce_fwd_impl(logits, target, target_dtype, n_img, num_classes, plane, ignore_index, with_rce, out, workspace,
                           stream, tss::LossMix{nullptr, 0.f, 0.f, 0, 0});
}

extern "C" int tss_ce_rce_fwd_mix(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                                  int64_t plane, int ignore_index, int with_rce, double alpha, double beta, int ce_sum,
                                  double* out, float* losses, void* workspace, void* stream) {
    TSS_REQUIRE(losses != nullptr, TSS_E_ARG, "tss_ce_rce_fwd_mix: null losses");
    return ce_rce_fwd_impl(logits, target, target_dtype, n_img, num_classes, plane, ignore_index, with_rce, out, workspace,
                           stream, tss::LossMix{losses, static_cast<float>(alpha), static_cast<float>(beta), with_rce, ce_sum});
}

static int ce_rce_fwd_impl(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                           int64_t plane, int ignore_index, int with_rce, double* out, void* workspace, void* stream,
                           const tss::LossMix mix) {
    using namespace tss;
    if (int r = check_loss_args("tss_ce_rce_fwd", logits, target, target_dtype, n_img, num_classes, plane)) return r;
    TSS_REQUIRE(out && workspace, TSS_E_ARG, "tss_ce_rce_fwd: null out/workspace");
    cudaStream_t st = as_stream(stream);
    const bool vec = plane % 4 == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0;
    // vectorised form: exactly the resident CTAs (2 per SM at 120 registers) in a grid-stride loop -- ~16 quads per thread, so
    // the last partial iteration is a 5 % tail; with 8 CTAs per SM it was 4.05 quads per thread: a fifth iteration for a few
    // per cent of the CTAs = 20 % tail.  (Contiguous per-CTA ranges were slower: 78 vs 67 us.)
    int grid = loss_grid(vec ? n_img * plane / 4 : n_img * plane);
    if (vec && grid > kNumSMs * TSS_LOSS_FWD_CTAS_PER_SM) grid = kNumSMs * TSS_LOSS_FWD_CTAS_PER_SM;
    double* part = static_cast<double*>(workspace);
    unsigned int* ticket = ticket_of(workspace);
    if (vec) TSS_DISPATCH_CT_TV(ce_rce_fwd_kernel, true, num_classes, target_dtype, grid, st, logits, target, n_img, num_classes, plane,
                                ignore_index, with_rce, part, ticket, out, mix);
    else TSS_DISPATCH_CT_TV(ce_rce_fwd_kernel, false, num_classes, target_dtype, grid, st, logits, target, n_img, num_classes, plane,
                            ignore_index, with_rce, part, ticket, out, mix);
    TSS_LAUNCH_CHECK("tss_ce_rce_fwd");
    return TSS_OK;
}

static int ce_rce_bwd_impl(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                           int64_t plane, int ignore_index, const float* scales, float* grad, void* stream,
                           const tss::GradMix gm) {
    using namespace tss;
    if (int r = check_loss_args("tss_ce_rce_bwd", logits, target, target_dtype, n_img, num_classes, plane)) return r;
    TSS_REQUIRE((scales || gm.grads) && grad, TSS_E_ARG, "tss_ce_rce_bwd: null scales/grad");
    if (n_img == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    const int grid = loss_grid(n_img * plane);
    TSS_DISPATCH_CT_T(ce_rce_bwd_kernel, num_classes, target_dtype, grid, st, logits, target, n_img, num_classes, plane,
                      ignore_index, scales, grad, gm);
    TSS_LAUNCH_CHECK("tss_ce_rce_bwd");
    return TSS_OK;
}

extern "C" int tss_ce_rce_bwd_mix(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                                  int64_t plane, int ignore_index, int with_rce, double alpha, double beta, int ce_sum,
                                  const float* upstream3, const double* fwd_out, float* grad, void* stream) {
    TSS_REQUIRE(upstream3 && fwd_out, TSS_E_ARG, "tss_ce_rce_bwd_mix: null upstream gradients / forward sums");
    return ce_rce_bwd_impl(logits, target, target_dtype, n_img, num_classes, plane, ignore_index, nullptr, grad, stream,
                           tss::GradMix{upstream3, fwd_out, static_cast<float>(alpha), static_cast<float>(beta), with_rce, ce_sum});
}

extern "C" int tss_ce_rce_bwd(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                              int64_t plane, int ignore_index, const float* scales, float* grad, void* stream) {
    return ce_rce_bwd_impl(logits, target, target_dtype, n_img, num_classes, plane, ignore_index, scales, grad, stream,
                           tss::GradMix{nullptr, nullptr, 0.f, 0.f, 0, 0});
}

extern "C" int tss_focal_fwd(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                             int64_t plane, float gamma, const float* alpha, double* out, void* workspace, void* stream) {
    using namespace tss;
    if (int r = check_loss_args("tss_focal_fwd", logits, target, target_dtype, n_img, num_classes, plane)) return r;
    TSS_REQUIRE(out && workspace, TSS_E_ARG, "tss_focal_fwd: null out/workspace");
    cudaStream_t st = as_stream(stream);
    const int grid = loss_grid(n_img * plane);
    double* part = static_cast<double*>(workspace);
    TSS_DISPATCH_CT_T(focal_fwd_kernel, num_classes, target_dtype, grid, st, logits, target, n_img, num_classes, plane, gamma,
                      alpha, part, ticket_of(workspace), out);
    TSS_LAUNCH_CHECK("tss_focal_fwd");
    return TSS_OK;
}

extern "C" int tss_focal_bwd(const float* logits, const void* target, int target_dtype, int64_t n_img, int num_classes,
                             int64_t plane, float gamma, const float* alpha, const float* scale, float* grad, void* stream) {
    using namespace tss;
    if (int r = check_loss_args("tss_focal_bwd", logits, target, target_dtype, n_img, num_classes, plane)) return r;
    TSS_REQUIRE(scale && grad, TSS_E_ARG, "tss_focal_bwd: null scale/grad");
    if (n_img == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    const int grid = loss_grid(n_img * plane);
    TSS_DISPATCH_CT_T(focal_bwd_kernel, num_classes, target_dtype, grid, st, logits, target, n_img, num_classes, plane, gamma,
                      alpha, scale, grad);
    TSS_LAUNCH_CHECK("tss_focal_bwd");
    return TSS_OK;
}
