// A1-A4: fused bilinear upsample -> per-pixel softmax / argmax / confidence -> fp16 + uint8 storage
// casts -> class-wise threshold into ignore_index (generate_pseudo_label.py:74-94,
// data/freiburg_dataset.py:115-123, segmentation_evaluate.py:38-42).
//
// pl_tile_kernel: one CTA = one 128-wide x TY-high tile of one image.
//   phase 1  horizontal taps: for every source row the tile touches (NR rows) and every class, the
//            x-interpolated value of each of the tile's 128 output columns goes to shared memory
//            (lx0*a + lx1*b: the reference's operation order, horizontal first);
//   phase 2  every thread owns 4 consecutive output pixels of a row: two 128-bit shared loads per class
//            give the two vertical taps, the C interpolated logits stay in registers, then max/argmax,
//            C ex2 on the SFU, reciprocal, storage casts and threshold; one 32-bit label store and one
//            64-bit fp16 store per thread (a warp writes 128 B / 256 B contiguous).
//   The C*H*W upsampled tensor the reference materialises (1 GB at 16x13x640x1920) never exists.
// pl_direct_kernel: same pixel reduction on full-resolution input (H==h, W==w), 128-bit coalesced
//   loads of the C planes -- the HBM-bound variant (4C+3 B/px).
#include "common.cuh"

namespace tss {

constexpr int kTX = 128;          // tile width: 32 lanes x 4 px
constexpr int kPLThreads = 256;   // 8 warps, each walks rows wy, wy+8, ...

struct PLArgs {
    const float* src;
    const float* src2;
    uint8_t* label;
    __half* conf16;
    float* conf32;
    uint8_t* label_thr;
    const uint8_t* gt;             // fused confusion matrix (register-resident kernel only): ground truth [N,H,W] uint8 ...
    unsigned long long* hist;      // ... and the [C,C] int64 matrix it accumulates into (rows = ground truth)
    int C, h, w, H, W;
    float sy, sx;
    int mode, flip;
    int TY, NR;
    int vec_ok;
    int has_thr, ignore;
    unsigned int keep_mask;
    float thr16[TSS_MAX_CLASSES];  // thresholds already rounded to fp16 (numpy compares in fp16)
    // one-lookup form of the threshold block: +inf for classes outside the whitelist (always rejected),
    // -inf when no threshold is given (never rejected), else thr16
    float thr_eff[TSS_MAX_CLASSES];
};

struct Tap {
    int i0, i1;
    float l0, l1;
};
// ATen area_pixel_compute_source_index, align_corners=True: src = scale * dst, all fp32
__device__ __forceinline__ Tap make_tap(float scale, int dst, int n_in) {
    Tap t;
    const float s = __fmul_rn(scale, static_cast<float>(dst));  // rounded product: never contracted into the subtraction
    t.i0 = min(static_cast<int>(s), n_in - 1);
    t.i1 = t.i0 + (t.i0 < n_in - 1 ? 1 : 0);
    t.l1 = __fsub_rn(s, static_cast<float>(t.i0));
    t.l0 = __fsub_rn(1.0f, t.l1);
    return t;
}

// per-pixel reduction over classes; v[] = the C values of one pixel
template <int CT>
__device__ __forceinline__ void reduce_pixel(const float (&v)[CT], int C, int mode, float& conf, int& label) {
    float m = v[0];
    int idx = 0;
#pragma unroll
    for (int c = 1; c < CT; ++c) {
        if (c < C) {
            const bool gt = v[c] > m;  // strict: ties keep the lowest class, as torch.max / argmax
            m = gt ? v[c] : m;
            idx = gt ? c : idx;
        }
    }
    label = idx;
    if (mode == TSS_PL_UP_SOFTMAX || mode == TSS_PL_PROB_UP_SOFTMAX) {
        const float mb = m * kLog2e;
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < C) s += ex2_approx(fmaf(v[c], kLog2e, -mb));
        conf = __fdiv_rn(1.0f, s);
    } else {
        conf = m;
    }
}

__device__ __forceinline__ uint8_t apply_threshold(const PLArgs& a, int label, float conf) {
    // fp16(conf) <= fp16(thr[label])  ->  ignore; classes outside the whitelist -> ignore
    const float c16 = __half2float(__float2half_rn(conf));
    const bool reject = a.has_thr && (c16 <= a.thr16[label]);
    const bool keep = (a.keep_mask >> label) & 1u;
    return static_cast<uint8_t>((reject || !keep) ? a.ignore : label);
}

__device__ __forceinline__ void store_quad(const PLArgs& a, int64_t off, int valid, const float (&conf)[4],
                                           const int (&lab)[4]) {
    if (a.vec_ok && valid == 4) {
        if (a.label)
            *reinterpret_cast<uint32_t*>(a.label + off) =
                static_cast<uint32_t>(lab[0]) | (static_cast<uint32_t>(lab[1]) << 8) |
                (static_cast<uint32_t>(lab[2]) << 16) | (static_cast<uint32_t>(lab[3]) << 24);
        if (a.conf16) {
            __half2 lo = __floats2half2_rn(conf[0], conf[1]);
            __half2 hi = __floats2half2_rn(conf[2], conf[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(a.conf16 + off) = pk;
        }
        if (a.conf32) *reinterpret_cast<float4*>(a.conf32 + off) = make_float4(conf[0], conf[1], conf[2], conf[3]);
        if (a.label_thr) {
            uint32_t t = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) t |= static_cast<uint32_t>(apply_threshold(a, lab[i], conf[i])) << (8 * i);
            *reinterpret_cast<uint32_t*>(a.label_thr + off) = t;
        }
    } else {
        for (int i = 0; i < valid; ++i) {
            if (a.label) a.label[off + i] = static_cast<uint8_t>(lab[i]);
            if (a.conf16) a.conf16[off + i] = __float2half_rn(conf[i]);
            if (a.conf32) a.conf32[off + i] = conf[i];
            if (a.label_thr) a.label_thr[off + i] = apply_threshold(a, lab[i], conf[i]);
        }
    }
}

template <int CT, bool FLIP>
__global__ void __launch_bounds__(kPLThreads, 2) pl_tile_kernel(const PLArgs a) {
    extern __shared__ __align__(16) float hs[];  // [FLIP ? 2 : 1][C][NR][kTX]
    const int C = a.C, NR = a.NR;
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * a.TY;
    const int64_t n = blockIdx.z;
    const int rb = min(static_cast<int>(a.sy * static_cast<float>(y0)), a.h - 1);
    const int plane_sz = C * NR * kTX;

    // ---- phase 1: horizontal interpolation of the NR source rows this tile needs
    {
        const int tx = tid & (kTX - 1);
        const int gx = min(x0 + tx, a.W - 1);
        const Tap t = make_tap(a.sx, gx, a.w);
        Tap t2 = t;
        if (FLIP && a.flip == TSS_FLIP_WIDTH) t2 = make_tap(a.sx, a.W - 1 - gx, a.w);
        const float* base = a.src + n * C * static_cast<int64_t>(a.h) * a.w;
        const float* base2 = FLIP ? a.src2 + n * C * static_cast<int64_t>(a.h) * a.w : nullptr;
        for (int j = tid / kTX; j < C * NR; j += kPLThreads / kTX) {
            const int c = j / NR, r = j - c * NR;
            const int row = min(rb + r, a.h - 1);
            const float* p = base + (static_cast<int64_t>(c) * a.h + row) * a.w;
            hs[j * kTX + tx] = fmaf(t.l1, __ldg(p + t.i1), t.l0 * __ldg(p + t.i0));
            if (FLIP) {
                const int c2 = (a.flip == TSS_FLIP_CLASS) ? (C - 1 - c) : c;
                const float* q = base2 + (static_cast<int64_t>(c2) * a.h + row) * a.w;
                hs[plane_sz + j * kTX + tx] = fmaf(t2.l1, __ldg(q + t2.i1), t2.l0 * __ldg(q + t2.i0));
            }
        }
    }
    __syncthreads();

    // ---- phase 2: vertical interpolation + per-pixel reduction, 4 px per thread
    const int warp = tid >> 5, lane = tid & 31;
    const int xq = x0 + 4 * lane;
    if (xq >= a.W) return;
    const int valid = min(4, a.W - xq);
    for (int ry = warp; ry < a.TY; ry += kPLThreads / 32) {
        const int y = y0 + ry;
        if (y >= a.H) break;
        const Tap ty = make_tap(a.sy, y, a.h);
        const int r0 = ty.i0 - rb, r1 = ty.i1 - rb;
        float v[4][CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) {
            if (c < C) {
                const float4 t0 = *reinterpret_cast<const float4*>(hs + (c * NR + r0) * kTX + 4 * lane);
                const float4 t1 = *reinterpret_cast<const float4*>(hs + (c * NR + r1) * kTX + 4 * lane);
                v[0][c] = fmaf(ty.l1, t1.x, ty.l0 * t0.x);
                v[1][c] = fmaf(ty.l1, t1.y, ty.l0 * t0.y);
                v[2][c] = fmaf(ty.l1, t1.z, ty.l0 * t0.z);
                v[3][c] = fmaf(ty.l1, t1.w, ty.l0 * t0.w);
                if (FLIP) {
                    const float4 u0 = *reinterpret_cast<const float4*>(hs + plane_sz + (c * NR + r0) * kTX + 4 * lane);
                    const float4 u1 = *reinterpret_cast<const float4*>(hs + plane_sz + (c * NR + r1) * kTX + 4 * lane);
                    v[0][c] = (v[0][c] + fmaf(ty.l1, u1.x, ty.l0 * u0.x)) * 0.5f;
                    v[1][c] = (v[1][c] + fmaf(ty.l1, u1.y, ty.l0 * u0.y)) * 0.5f;
                    v[2][c] = (v[2][c] + fmaf(ty.l1, u1.z, ty.l0 * u0.z)) * 0.5f;
                    v[3][c] = (v[3][c] + fmaf(ty.l1, u1.w, ty.l0 * u0.w)) * 0.5f;
                }
            }
        }
        float conf[4];
        int lab[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) reduce_pixel<CT>(v[i], C, a.mode, conf[i], lab[i]);
        store_quad(a, (n * a.H + y) * static_cast<int64_t>(a.W) + xq, valid, conf, lab);
    }
}

// ------------------------------------------------------------------------------------------------
// pl_up_kernel: the headline path (logits -> upsample -> softmax/argmax -> casts -> threshold) without
// shared memory.  A warp owns a 64-pixel-wide column block and a contiguous range of output rows of the
// flattened [N*H] row space; every lane owns TWO adjacent pixels (packed f32x2 arithmetic) and keeps the
// horizontally interpolated values of the two source rows that bracket the current output row in
// registers (2 rows x C classes x 2 px).  Walking down, a new source row is needed once per ~scale output
// rows: 4 read-only loads per class (neighbouring lanes hit the same L1 lines -- the 16 MB low-res logits
// are read from HBM once).  Per output pixel that leaves the vertical lerp, the class reduction and the
// C exponentials -- the SFU (MUFU.EX2, 16 lanes/clk/SM) is the bound, not HBM.
// Bit-identical to pl_tile_kernel (same operation order).
constexpr int kUpBlockPx = 64;

// value at base[off] with the address formed by ONE 32x32+64 multiply-add (off = plane/row offset in
// elements, uniform across the CTA; base = the lane's own tap pointer)
__device__ __forceinline__ float ldg_at(const float* base, unsigned off) {
    const float* p;
    asm("mad.wide.u32 %0, %1, 4, %2;" : "=l"(p) : "r"(off), "l"(base));
    return __ldg(p);
}

template <int C>
__device__ __forceinline__ void up_load_row(f32x2 (&dst)[C], const float* pa0, const float* pa1, const float* pb0,
                                            const float* pb1, unsigned hw, unsigned row_off, f32x2 lx0, f32x2 lx1) {
    unsigned off = row_off;  // one image's C planes stay below 2^30 elements (checked on the host)
#pragma unroll
    for (int c = 0; c < C; ++c, off += hw) {
        const f32x2 lo = pack2(ldg_at(pa0, off), ldg_at(pb0, off));
        const f32x2 hi = pack2(ldg_at(pa1, off), ldg_at(pb1, off));
        dst[c] = fma2(lx1, hi, mul2(lx0, lo));  // lx1*b + (lx0*a), as the tile kernel and ATen order it
    }
}

// Cooperative form for upsampling factors >= ~2.1: the 64 output columns of the warp touch at most 32
// source columns, so ONE coalesced load per class fetches them (lane j holds column base+j) and every lane
// picks its four taps with shuffles -- 1 load + 4 SHFL per class instead of 4 loads with 64-bit addresses.
template <int C>
__device__ __forceinline__ void up_load_row_coop(f32x2 (&dst)[C], const float* pcol, unsigned hw, unsigned row_off,
                                                 int ra0, int ra1, int rb0, int rb1, f32x2 lx0, f32x2 lx1) {
    unsigned off = row_off;
    float x[C];
#pragma unroll
    for (int c = 0; c < C; ++c, off += hw) x[c] = ldg_at(pcol, off);
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const f32x2 lo = pack2(__shfl_sync(0xffffffffu, x[c], ra0), __shfl_sync(0xffffffffu, x[c], rb0));
        const f32x2 hi = pack2(__shfl_sync(0xffffffffu, x[c], ra1), __shfl_sync(0xffffffffu, x[c], rb1));
        dst[c] = fma2(lx1, hi, mul2(lx0, lo));
    }
}

__device__ __forceinline__ float set_neu(float a, float b) {  // (a != b or unordered) ? 1.0f : 0.0f  (FSET.BF)
    float d;
    asm("set.neu.f32.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}

__device__ __forceinline__ float rcp_refined(float s) {
    // s in [1, C]: MUFU.RCP (<= 1 ulp) + one Newton step; no special-case path
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(s));
    return fmaf(r, fmaf(-s, r, 1.0f), r);
}

// epilogue of one finished row: confidence, storage casts, threshold, stores.  OUTS = compile-time set of
// outputs (bit 0 label, 1 conf16, 2 conf32, 3 label_thr) or -1 for run-time pointer tests; `off` = element
// offset g*W + x of the pixel pair in every [N,H,W] output.
constexpr int kOutsRuntime = -1;
constexpr int kOutsPseudoLabel = 0b1011;  // label + conf16 + label_thr: the pseudo-label stage
template <int MODE, int OUTS>
__device__ __forceinline__ void up_emit(const PLArgs& a, bool valid, int64_t off, float sa, float sb, int ia, int ib) {
    float ca = sa, cb = sb;
    if (MODE == TSS_PL_UP_SOFTMAX) {
        ca = rcp_refined(sa);
        cb = rcp_refined(sb);
    }
    const bool w_label = OUTS == kOutsRuntime ? a.label != nullptr : (OUTS & 1) != 0;
    const bool w_c16 = OUTS == kOutsRuntime ? a.conf16 != nullptr : (OUTS & 2) != 0;
    const bool w_c32 = OUTS == kOutsRuntime ? a.conf32 != nullptr : (OUTS & 4) != 0;
    const bool w_thr = OUTS == kOutsRuntime ? a.label_thr != nullptr : (OUTS & 8) != 0;
    const __half2 c16 = __floats2half2_rn(ca, cb);
    if (valid && w_label) *reinterpret_cast<uint16_t*>(a.label + off) = static_cast<uint16_t>(ia | (ib << 8));
    if (valid && w_c16) *reinterpret_cast<__half2*>(a.conf16 + off) = c16;
    if (valid && w_c32) *reinterpret_cast<float2*>(a.conf32 + off) = make_float2(ca, cb);
    if (w_thr) {
        const float2 cf = __half22float2(c16);
        const int ta = (cf.x <= a.thr_eff[ia]) ? a.ignore : ia;
        const int tb = (cf.y <= a.thr_eff[ib]) ? a.ignore : ib;
        if (valid) *reinterpret_cast<uint16_t*>(a.label_thr + off) = static_cast<uint16_t>(ta | (tb << 8));
    }
}

template <int C>
struct UpCfg {  // register budget: 96 (20 warps/SM) holds C <= 13; 19 classes need the 128-register cap (16 warps/SM)
#ifndef TSS_UP_PIPE
#define TSS_UP_PIPE 0
#endif
#ifndef TSS_UP_WARPS13
#define TSS_UP_WARPS13 16
#endif
    static constexpr int kWarpsPerSM = C <= 13 ? TSS_UP_WARPS13 : 16;
};

// the per-row arithmetic of pl_up_kernel for NR output rows that share one bracket of source rows (A, B): every stage
// loops over the NR rows inside its class loop, so that the rows' dependency chains (max tree, argmax Horner chain,
// exponent sums) are interleaved in one instruction stream -- ILP for a kernel whose largest stall is the fixed-latency
// dependency wait
template <int C, int MODE, int OUTS, int NR>
__device__ __forceinline__ void up_rows(const PLArgs& a, const f32x2 (&A)[C], const f32x2 (&B)[C], const float (&l0)[NR],
                                        const float (&l1)[NR], bool active, int64_t off) {
    float va[NR][C], vb[NR][C];
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int j = 0; j < NR; ++j)
            unpack2(fma2(pack2(l1[j], l1[j]), B[c], mul2(pack2(l0[j], l0[j]), A[c])), va[j][c], vb[j][c]);
    float ma[NR], mb[NR];
#pragma unroll
    for (int j = 0; j < NR; ++j) { ma[j] = va[j][0]; mb[j] = vb[j][0]; }
#pragma unroll
    for (int c = 1; c < C; ++c)
#pragma unroll
        for (int j = 0; j < NR; ++j) { ma[j] = fmaxf(ma[j], va[j][c]); mb[j] = fmaxf(mb[j], vb[j][c]); }
    f32x2 hidx[NR];
#pragma unroll
    for (int j = 0; j < NR; ++j) hidx[j] = pack2(0.f, 0.f);
#pragma unroll
    for (int c = C - 2; c >= 0; --c)
#pragma unroll
        for (int j = 0; j < NR; ++j) {
            const f32x2 ne = pack2(set_neu(va[j][c], ma[j]), set_neu(vb[j][c], mb[j]));
            hidx[j] = fma2(ne, hidx[j], ne);
        }
    float sa[NR], sb[NR];
    if (MODE == TSS_PL_UP_SOFTMAX) {
        const f32x2 l2e = pack2(kLog2e, kLog2e);
        f32x2 nmb[NR], acc[NR][4];
#pragma unroll
        for (int j = 0; j < NR; ++j) nmb[j] = pack2(-(ma[j] * kLog2e), -(mb[j] * kLog2e));
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int j = 0; j < NR; ++j) {
                float ta_, tb_;
                unpack2(fma2(pack2(va[j][c], vb[j][c]), l2e, nmb[j]), ta_, tb_);
                const f32x2 e = pack2(ex2_approx(ta_), ex2_approx(tb_));
                acc[j][c & 3] = c < 4 ? e : add2(acc[j][c & 3], e);
            }
#pragma unroll
        for (int j = 0; j < NR; ++j) unpack2(add2(add2(acc[j][0], acc[j][1]), add2(acc[j][2], acc[j][3])), sa[j], sb[j]);
    } else {
#pragma unroll
        for (int j = 0; j < NR; ++j) { sa[j] = ma[j]; sb[j] = mb[j]; }
    }
#pragma unroll
    for (int j = 0; j < NR; ++j) {
        float fia, fib;
        unpack2(add2(hidx[j], pack2(8388608.f, 8388608.f)), fia, fib);  // integer in the low mantissa bits
        up_emit<MODE, OUTS>(a, active, off + j * static_cast<int64_t>(a.W), sa[j], sb[j], __float_as_int(fia) & 0xff,
                            __float_as_int(fib) & 0xff);
    }
}

// RC = register class: 0 = the 128-register cap (16 warps per SM alone), 1 = a 96-register cap (20 warps per SM; a few
// spilled words at C = 13): TSS_PL_FLAG_REGS96, for the co-resident forms where the other kernel's CTA leaves only
// about 43 K registers per SM -- 13 warps of this class instead of 10.
// HIST: the confusion matrix of (ground truth, label) is accumulated in the same pass (utils/eval_tools.py:4-9 on the
// labels of generate_pseudo_label.py:85 / segmentation_evaluate.py:42): one 16-bit load of the pixel pair's ground
// truth per row and two shared-memory atomics on a C x C int32 matrix per CTA -- LSU work in a kernel bound by the XU
// pipe and the issue slots -- flushed with one 64-bit global atomic per non-zero bin when the CTA ends.  The separate
// histogram kernel re-reads labels and ground truth (2 B/px) and takes 23 us at cfg2.
template <int C, int MODE, bool COOP, int OUTS, int NR, int RC = 0, bool HIST = false>
__global__ void __launch_bounds__(NR == 2 ? 384 : (RC == 1 ? 640 : UpCfg<C>::kWarpsPerSM * 32), 1)
pl_up_kernel(const PLArgs a, int64_t total_rows, int rows_per_seg, int col_blocks, int tab_rows) {
    // blockIdx.y = row segment, so the image / row bookkeeping and every row base address are uniform
    // across the CTA (uniform datapath); warps of a CTA take neighbouring column blocks.
    const int lane = threadIdx.x & 31;
    const int cb = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t g0 = static_cast<int64_t>(blockIdx.y) * rows_per_seg;
    const int64_t g1 = min(g0 + rows_per_seg, total_rows);
    // The vertical taps depend on the output row alone and every warp of the CTA walks the same rows: they are
    // computed once per CTA into shared memory (one broadcast 128-bit load per row instead of two int<->float
    // conversions on the XU pipe -- the busiest pipe of this kernel -- and eight more instructions per warp and row).
    extern __shared__ __align__(16) int4 ytab[];
    unsigned int* sh_hist = reinterpret_cast<unsigned int*>(ytab + tab_rows);  // HIST: C*C bins behind the tap table
    if (tab_rows > 0) {
        for (int r = threadIdx.x; r < static_cast<int>(g1 - g0); r += blockDim.x) {
            const Tap t = make_tap(a.sy, static_cast<int>((g0 + r) % a.H), a.h);
            ytab[r] = make_int4(t.i0, t.i1, __float_as_int(t.l0), __float_as_int(t.l1));
        }
    }
    if (HIST)
        for (int i = threadIdx.x; i < C * C; i += blockDim.x) sh_hist[i] = 0u;
    if (tab_rows > 0 || HIST) __syncthreads();
    if (!HIST && cb >= col_blocks) return;   // (with HIST every warp stays for the flush at the end)
    const int gx = cb * kUpBlockPx + 2 * lane;
    const bool active = gx < a.W;              // W is even on this path: both pixels valid or neither
    const int gxc = active ? gx : a.W - 2;
    const Tap ta = make_tap(a.sx, gxc, a.w), tb = make_tap(a.sx, gxc + 1, a.w);
    const f32x2 lx0 = pack2(ta.l0, tb.l0), lx1 = pack2(ta.l1, tb.l1);
    const int64_t img_stride = static_cast<int64_t>(C) * a.h * a.w;
    const unsigned hw = static_cast<unsigned>(a.h) * a.w;
    // cooperative load: lane j holds source column base + j of the current row / class
    const int base_col = make_tap(a.sx, min(cb * kUpBlockPx, a.W - 2), a.w).i0;
    const int my_col = min(base_col + lane, a.w - 1);
    const int ra0 = ta.i0 - base_col, ra1 = ta.i1 - base_col, rb0 = tb.i0 - base_col, rb1 = tb.i1 - base_col;

    const float *pa0 = nullptr, *pa1 = nullptr, *pb0 = nullptr, *pb1 = nullptr, *pcol = nullptr;
    f32x2 A[C], B[C];
    int cur0 = -1, cur1 = -1;
    int64_t n = g0 / a.H;
    int y = static_cast<int>(g0 - n * a.H);
    bool fresh = true;  // image changed (or first row): both bracketing rows must be loaded
    bool p_valid = false;  // a finished row waits for its epilogue
    int64_t p_off = 0;
    float p_sa = 1.f, p_sb = 1.f;
    int p_ia = 0, p_ib = 0;
    int64_t off = g0 * a.W + gx;
    const int n_rows = (HIST && cb >= col_blocks) ? 0 : static_cast<int>(g1 - g0);
    for (int r = 0; r < n_rows; ++r) {
        Tap ty;
        if (tab_rows > 0) {
            const int4 t = ytab[r];
            ty.i0 = t.x; ty.i1 = t.y; ty.l0 = __int_as_float(t.z); ty.l1 = __int_as_float(t.w);
        } else {
            ty = make_tap(a.sy, y, a.h);
        }
        if (fresh || ty.i0 != cur0 || ty.i1 != cur1) {
            if (fresh) {
                const float* img = a.src + n * img_stride;
                if (COOP) {
                    pcol = img + my_col;
                } else {
                    pa0 = img + ta.i0; pa1 = img + ta.i1; pb0 = img + tb.i0; pb1 = img + tb.i1;
                }
            }
            if (!fresh && ty.i0 == cur1) {
#pragma unroll
                for (int c = 0; c < C; ++c) A[c] = B[c];
            } else if (COOP) {
                up_load_row_coop<C>(A, pcol, hw, static_cast<unsigned>(ty.i0 * a.w), ra0, ra1, rb0, rb1, lx0, lx1);
            } else {
                up_load_row<C>(A, pa0, pa1, pb0, pb1, hw, static_cast<unsigned>(ty.i0 * a.w), lx0, lx1);
            }
            if (ty.i1 == ty.i0) {
#pragma unroll
                for (int c = 0; c < C; ++c) B[c] = A[c];
            } else if (COOP) {
                up_load_row_coop<C>(B, pcol, hw, static_cast<unsigned>(ty.i1 * a.w), ra0, ra1, rb0, rb1, lx0, lx1);
            } else {
                up_load_row<C>(B, pa0, pa1, pb0, pb1, hw, static_cast<unsigned>(ty.i1 * a.w), lx0, lx1);
            }
            cur0 = ty.i0;
            cur1 = ty.i1;
            fresh = false;
        }
        unsigned gtp = 0xffffu;   // ground truth of the pixel pair: issued here, consumed after the row's arithmetic
        if (HIST && active) gtp = *reinterpret_cast<const uint16_t*>(a.gt + off);
        if (NR == 2 && tab_rows > 0 && r + 1 < n_rows && y + 1 < a.H) {
            // the next output row lies between the same two source rows: both rows in one instruction stream
            const int4 t1 = ytab[r + 1];
            if (t1.x == cur0 && t1.y == cur1) {
                const float l0[2] = {ty.l0, __int_as_float(t1.z)}, l1[2] = {ty.l1, __int_as_float(t1.w)};
                up_rows<C, MODE, OUTS, 2>(a, A, B, l0, l1, active, off);
                off += 2 * static_cast<int64_t>(a.W);
                y += 2;
                ++r;
                if (y == a.H) { y = 0; ++n; fresh = true; }
                continue;
            }
        }
        // vertical lerp, two pixels per instruction
        const f32x2 ly0 = pack2(ty.l0, ty.l0), ly1 = pack2(ty.l1, ty.l1);
        float va[C], vb[C];
#pragma unroll
        for (int c = 0; c < C; ++c) unpack2(fma2(ly1, B[c], mul2(ly0, A[c])), va[c], vb[c]);
        float ma = va[0], mb = vb[0];
#pragma unroll
        for (int c = 1; c < C; ++c) { ma = fmaxf(ma, va[c]); mb = fmaxf(mb, vb[c]); }
        // lowest class equal to the maximum (torch.max tie rule) = the number of leading classes that differ
        // from it: with ne_c = (v_c != max) as 1.0 / 0.0, idx = ne_0 (1 + ne_1 (1 + ne_2 (...))) -- one FSET per
        // class and pixel plus ONE packed FFMA per class for the pixel pair, instead of compare + select chains
        // on the (half-rate) ALU pipe.  Exact: every partial value is an integer <= C.
        f32x2 hidx = pack2(0.f, 0.f);
#ifndef TSS_PL_EXP
#define TSS_PL_EXP 0   // experiment builds only (tools/pl_experiments.sh): 1 = no argmax chain, 2 = no exponentials, 4 = no stores
#endif
#if !(TSS_PL_EXP & 1)
#pragma unroll
        for (int c = C - 2; c >= 0; --c) {
            const f32x2 ne = pack2(set_neu(va[c], ma), set_neu(vb[c], mb));
            hidx = fma2(ne, hidx, ne);
        }
#endif
        float fia, fib;
        unpack2(add2(hidx, pack2(8388608.f, 8388608.f)), fia, fib);  // integer in the low mantissa bits
        const int ia = __float_as_int(fia) & 0xff, ib = __float_as_int(fib) & 0xff;
        float sa = ma, sb = mb;  // UP_ARGMAX: the "confidence" is the max logit
        if (MODE == TSS_PL_UP_SOFTMAX && !(TSS_PL_EXP & 2)) {
            const f32x2 l2e = pack2(kLog2e, kLog2e);
            const f32x2 nmb = pack2(-(ma * kLog2e), -(mb * kLog2e));
            f32x2 acc[4];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                float ta_, tb_;
                unpack2(fma2(pack2(va[c], vb[c]), l2e, nmb), ta_, tb_);
                const f32x2 e = pack2(ex2_approx(ta_), ex2_approx(tb_));
                acc[c & 3] = c < 4 ? e : add2(acc[c & 3], e);
            }
            unpack2(add2(add2(acc[0], acc[1]), add2(acc[2], acc[3])), sa, sb);
        }
        // Software pipeline: the reciprocal / casts / threshold / stores of THIS row are issued while the
        // next row's lerps and exponentials run (they are independent instruction streams in one block).
#if TSS_UP_PIPE
        up_emit<MODE, OUTS>(a, p_valid, p_off, p_sa, p_sb, p_ia, p_ib);
        p_valid = active;
        p_off = off;
        p_sa = sa; p_sb = sb; p_ia = ia; p_ib = ib;
#else
        up_emit<MODE, OUTS>(a, (TSS_PL_EXP & 4) ? (active && sa == -123.f) : active, off, sa, sb, ia, ib);
#endif
        if (HIST) {   // a pixel counts iff its ground truth is a class (inactive lanes carry 0xff)
            const unsigned ga = gtp & 0xffu, gb = gtp >> 8;
            if (ga < static_cast<unsigned>(C)) atomicAdd(&sh_hist[ga * C + ia], 1u);
            if (gb < static_cast<unsigned>(C)) atomicAdd(&sh_hist[gb * C + ib], 1u);
        }
        off += a.W;  // rows of consecutive images are contiguous in every [N,H,W] output
        if (++y == a.H) { y = 0; ++n; fresh = true; }
    }
#if TSS_UP_PIPE
    up_emit<MODE, OUTS>(a, p_valid, p_off, p_sa, p_sb, p_ia, p_ib);
#endif
    if (HIST) {
        __syncthreads();
        for (int i = threadIdx.x; i < C * C; i += blockDim.x) {
            const unsigned int v = sh_hist[i];
            if (v) atomicAdd(&a.hist[i], static_cast<unsigned long long>(v));
        }
    }
}

// full-resolution input: grid-stride over pixel quads of the flattened [N, H*W] map
template <int CT>
__global__ void __launch_bounds__(256) pl_direct_kernel(const PLArgs a, int64_t n_img) {
    const int C = a.C;
    const int64_t plane = static_cast<int64_t>(a.H) * a.W;
    const int64_t quads_per_img = (plane + 3) / 4;
    const int64_t total = n_img * quads_per_img;
    const int64_t q0 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    const int64_t qs = static_cast<int64_t>(gridDim.x) * blockDim.x;
    PixelWalk w(q0, qs, quads_per_img);   // (image, quad) without a 64-bit division per iteration
    for (int64_t q = q0; q < total; q += qs, w.next()) {
        const int64_t n = w.n;
        const int64_t p = w.p * 4;
        const int valid = static_cast<int>(min(static_cast<int64_t>(4), plane - p));
        const float* base = a.src + n * C * plane + p;
        float v[4][CT];
        if (a.vec_ok && valid == 4) {
#pragma unroll
            for (int c = 0; c < CT; ++c) {  // unconditional (clamped class): all C 128-bit loads in flight together
                const float4 t = ldg_stream_f4(base + static_cast<int64_t>(min(c, C - 1)) * plane);
                v[0][c] = t.x; v[1][c] = t.y; v[2][c] = t.z; v[3][c] = t.w;
            }
        } else {
#pragma unroll
            for (int c = 0; c < CT; ++c)
                if (c < C)
                    for (int i = 0; i < 4; ++i) v[i][c] = (i < valid) ? base[c * plane + i] : 0.f;
        }
        float conf[4];
        int lab[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) reduce_pixel<CT>(v[i], C, a.mode, conf[i], lab[i]);
        store_quad(a, n * plane + p, valid, conf, lab);
    }
}

// ------------------------------------------------------------------------------------------------
// A2 low-resolution softmax over the class axis of [N, C, P]
template <int CT>
__global__ void __launch_bounds__(256) softmax_nchw_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                           int64_t n_img, int C, int64_t plane) {
    const int64_t total = n_img * plane;
    const int64_t i0 = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    const int64_t is = static_cast<int64_t>(gridDim.x) * blockDim.x;
    PixelWalk w(i0, is, plane);
    for (int64_t i = i0; i < total; i += is, w.next()) {
        const int64_t n = w.n, p = w.p;
        const float* px = x + n * C * plane + p;
        float v[CT];
#pragma unroll
        for (int c = 0; c < CT; ++c) v[c] = __ldg(px + static_cast<int64_t>(min(c, C - 1)) * plane);  // all in flight
        float m = -INFINITY;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < C) m = fmaxf(m, v[c]);
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < C) {
                v[c] = expf(v[c] - m);  // full-precision exp: these probabilities are a stored product
                s += v[c];
            }
        float* py = y + n * C * plane + p;
#pragma unroll
        for (int c = 0; c < CT; ++c)
            if (c < C) py[c * plane] = __fdiv_rn(v[c], s);
    }
}

// A1 plain bilinear upsample of [planes, h, w] -> [planes, H, W]
__global__ void __launch_bounds__(256) upsample_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                       int64_t planes, int h, int w, int H, int W, float sy, float sx) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const Tap tx = make_tap(sx, x, w), ty = make_tap(sy, y, h);
    for (int64_t pl = blockIdx.z; pl < planes; pl += gridDim.z) {
        const float* p = src + pl * static_cast<int64_t>(h) * w;
        const float top = fmaf(tx.l1, __ldg(p + ty.i0 * w + tx.i1), tx.l0 * __ldg(p + ty.i0 * w + tx.i0));
        const float bot = fmaf(tx.l1, __ldg(p + ty.i1 * w + tx.i1), tx.l0 * __ldg(p + ty.i1 * w + tx.i0));
        dst[(pl * H + y) * static_cast<int64_t>(W) + x] = fmaf(ty.l1, bot, ty.l0 * top);
    }
}

__global__ void __launch_bounds__(256) threshold_kernel(const uint8_t* __restrict__ lab, const __half* __restrict__ conf,
                                                        int64_t n, PLArgs a, uint8_t* __restrict__ out) {
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int l = lab[i];
        uint8_t r = static_cast<uint8_t>(a.ignore);
        if (l < a.C) {
            const bool reject = a.has_thr && (__half2float(conf[i]) <= a.thr16[l]);
            const bool keep = (a.keep_mask >> l) & 1u;
            r = static_cast<uint8_t>((reject || !keep) ? a.ignore : l);
        }
        out[i] = r;
    }
}

__global__ void __launch_bounds__(256) gather_nearest_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                             int64_t planes, int h, int w, int H, int W,
                                                             const int32_t* __restrict__ ytab, const int32_t* __restrict__ xtab) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    const int sx = xtab[x], sy = ytab[y];
    for (int64_t pl = blockIdx.z; pl < planes; pl += gridDim.z)
        dst[(pl * H + y) * static_cast<int64_t>(W) + x] = src[(pl * h + sy) * static_cast<int64_t>(w) + sx];
}

static float scale_of(int n_in, int n_out) {
    // ATen area_pixel_compute_scale<float>(align_corners=true)
    return n_out > 1 ? static_cast<float>(n_in - 1) / static_cast<float>(n_out - 1) : 0.0f;
}

static void fill_threshold(PLArgs& a, int C, const float* thr_host, const uint8_t* keep_host, int ignore_index) {
    a.has_thr = thr_host != nullptr;
    a.ignore = ignore_index;
    a.keep_mask = 0;
    for (int c = 0; c < TSS_MAX_CLASSES; ++c) {
        a.thr16[c] = (thr_host && c < C) ? __half2float(__float2half_rn(thr_host[c])) : 0.f;
        const bool kept = c < C && (keep_host == nullptr || keep_host[c]);
        if (kept) a.keep_mask |= 1u << c;
        a.thr_eff[c] = !kept ? INFINITY : (thr_host ? a.thr16[c] : -INFINITY);
    }
}

template <int CT>
static int launch_tile(const PLArgs& a, int64_t n_img, size_t smem, cudaStream_t st) {
    dim3 grid((a.W + kTX - 1) / kTX, (a.H + a.TY - 1) / a.TY, static_cast<unsigned>(n_img));
    cudaError_t e;
    if (a.flip != TSS_FLIP_NONE) {
        e = cudaFuncSetAttribute(pl_tile_kernel<CT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_status(e, "tss_pseudo_label: smem attribute");
        pl_tile_kernel<CT, true><<<grid, kPLThreads, smem, st>>>(a);
    } else {
        e = cudaFuncSetAttribute(pl_tile_kernel<CT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_status(e, "tss_pseudo_label: smem attribute");
        pl_tile_kernel<CT, false><<<grid, kPLThreads, smem, st>>>(a);
    }
    TSS_LAUNCH_CHECK("tss_pseudo_label(tile)");
    return TSS_OK;
}

template <int CT>
static int launch_direct(const PLArgs& a, int64_t n_img, cudaStream_t st) {
    const int64_t quads = n_img * ((static_cast<int64_t>(a.H) * a.W + 3) / 4);
    int64_t blocks = (quads + 255) / 256;
    const int64_t cap = static_cast<int64_t>(kNumSMs) * 16;
    if (blocks > cap) blocks = cap;
    pl_direct_kernel<CT><<<static_cast<int>(blocks), 256, 0, st>>>(a, n_img);
    TSS_LAUNCH_CHECK("tss_pseudo_label(direct)");
    return TSS_OK;
}

template <int C, int MODE>
static int launch_up(const PLArgs& a, int64_t n_img, int flags, cudaStream_t st) {
    // The kernel is compiled for a register cap that lets an SM hold UpCfg<C>::kWarpsPerSM warps.  Default: ONE wave,
    // with the warps per CTA that wastes the fewest column slots and keeps the SM fullest.
    // flags (co-resident forms, see include/tss_b200.h): sm_warps sizes the wave for fewer warps per SM, leaving
    // registers for a co-resident kernel (the accumulate kernel of the prototype stage, which is HBM-bound while this
    // one is XU / issue bound); cta_warps fixes the CTA size and waves > 1 cuts the rows into that many times more
    // segments than CTA slots, so that the block scheduler refills an SM with pseudo-label CTAs as the co-resident
    // kernel's CTAs retire (a row segment still spans dozens of output rows: its two start-up row loads stay < 1 %).
    const int sm_warps = (flags >> 8) & 0xff, cta_warps = (flags >> 16) & 0xf, waves = ((flags >> 20) & 0xf) ? ((flags >> 20) & 0xf) : 1;
    const int col_blocks = (a.W + kUpBlockPx - 1) / kUpBlockPx;
    int best_wpc = 8, best_cps = 2;
    double best_eff = -1.0;
    const bool rows2 = (flags & TSS_PL_FLAG_ROWS2) != 0;
    const bool regs96 = (flags & TSS_PL_FLAG_REGS96) != 0 && C <= 13 && !rows2;
    const int kWmax = rows2 ? 12 : (regs96 ? 20 : UpCfg<C>::kWarpsPerSM);  // the two-row form holds 170 registers per thread
    const int kW = (sm_warps >= 3 && sm_warps < kWmax) ? sm_warps : kWmax;
    for (int wpc = (cta_warps ? cta_warps : 4); wpc <= (cta_warps ? cta_warps : kW); ++wpc) {
        const int cps = kW / wpc;                                  // CTAs per SM by registers
        if (cps < 1) break;
        const int gx = (col_blocks + wpc - 1) / wpc;
        const double eff = (static_cast<double>(col_blocks) / (gx * wpc)) * (cps * wpc / static_cast<double>(kW));
        if (eff > best_eff + 1e-9) { best_eff = eff; best_wpc = wpc; best_cps = cps; }
    }
    const int gx = (col_blocks + best_wpc - 1) / best_wpc;
    const int64_t total_rows = n_img * a.H;
    int64_t segs = static_cast<int64_t>(kNumSMs) * best_cps * waves / gx;  // CTA slots (x waves) / CTAs per row segment
    if (segs < 1) segs = 1;
    int64_t rps = (total_rows + segs - 1) / segs;
    if (rps < 1) rps = 1;
    segs = (total_rows + rps - 1) / rps;
    TSS_REQUIRE(segs <= 65535, TSS_E_ARG, "tss_pseudo_label: too many row segments");
    dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(segs));
    // the 64 output columns of a warp span floor(sx*63)+2 source columns at most (+1 for fp32 slack)
    const bool coop = (a.sx * 63.0f + 3.0f) < 32.0f;
    const bool pl_outs = MODE == TSS_PL_UP_SOFTMAX && a.label && a.conf16 && !a.conf32 && a.label_thr;
    const int block = best_wpc * 32, r = static_cast<int>(rps);
    // Co-resident form: the accumulate kernel that shares the SMs needs the 228 KB shared-memory configuration, and an
    // SM cannot hold CTAs of two kernels that ask for different L1 / shared splits -- so this kernel (which uses no
    // shared memory) asks for the same split; on its own it keeps the default (largest L1).
    const int carve = (sm_warps > 0 || cta_warps > 0) ? static_cast<int>(cudaSharedmemCarveoutMaxShared) : static_cast<int>(cudaSharedmemCarveoutDefault);
    const int tab_rows = r <= 2048 ? r : 0;   // per-CTA vertical tap table (16 B per row of the segment), else computed per row
    const bool hist = a.hist != nullptr;
    auto go = [&](auto kern) -> cudaError_t {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        if (e != cudaSuccess) return e;
        const size_t smem = static_cast<size_t>(tab_rows) * sizeof(int4) + (hist ? static_cast<size_t>(C) * C * sizeof(unsigned int) : 0);
        kern<<<grid, block, smem, st>>>(a, total_rows, r, col_blocks, tab_rows);
        return cudaSuccess;
    };
    constexpr int kOutsFast = MODE == TSS_PL_UP_SOFTMAX ? kOutsPseudoLabel : kOutsRuntime;
    cudaError_t e;
    if (hist && pl_outs)          // (a histogram only comes here when the cooperative form applies: see pseudo_label_impl)
        e = go(pl_up_kernel<C, MODE, true, kOutsFast, 1, 0, true>);
    else if (hist)
        e = go(pl_up_kernel<C, MODE, true, kOutsRuntime, 1, 0, true>);
    else if (coop && pl_outs && rows2 && tab_rows > 0)
        e = go(pl_up_kernel<C, MODE, true, MODE == TSS_PL_UP_SOFTMAX ? kOutsPseudoLabel : kOutsRuntime, 2>);
    else if (coop && pl_outs && regs96)
        e = go(pl_up_kernel<C <= 13 ? C : 13, MODE, true, MODE == TSS_PL_UP_SOFTMAX ? kOutsPseudoLabel : kOutsRuntime, 1, 1>);
    else if (coop && pl_outs)
        e = go(pl_up_kernel<C, MODE, true, MODE == TSS_PL_UP_SOFTMAX ? kOutsPseudoLabel : kOutsRuntime, 1>);
    else if (coop)
        e = go(pl_up_kernel<C, MODE, true, kOutsRuntime, 1>);
    else
        e = go(pl_up_kernel<C, MODE, false, kOutsRuntime, 1>);
    if (e != cudaSuccess) return cuda_status(e, "tss_pseudo_label: carveout attribute");
    TSS_LAUNCH_CHECK("tss_pseudo_label(up)");
    return TSS_OK;
}

}  // namespace tss

namespace tss {
static int pseudo_label_impl(const float* src, const float* src_flip, int64_t n_img, int num_classes, int h, int w,
                             int H, int W, int mode, int flip, uint8_t* label_out, uint16_t* conf16_out,
                             float* conf32_out, uint8_t* label_thr_out, const float* thr_host,
                             const uint8_t* keep_host, int ignore_index, const uint8_t* gt, int64_t* hist, int flags,
                             void* stream) {
    const int g_pl_force_tile = (flags & TSS_PL_FLAG_FORCE_TILE) ? 1 : 0;
    const int sm_warps = (flags >> 8) & 0xff;
    TSS_REQUIRE((flags >> 24) == 0 && flags >= 0 && (flags & 0xf8) == 0, TSS_E_ARG, "tss_pseudo_label: bad flags 0x%x", flags);
    TSS_REQUIRE(src != nullptr, TSS_E_ARG, "tss_pseudo_label: null src");
    TSS_REQUIRE(n_img >= 0 && h > 0 && w > 0 && H > 0 && W > 0, TSS_E_ARG, "tss_pseudo_label: bad extents");
    TSS_REQUIRE(n_img <= 65535, TSS_E_ARG, "tss_pseudo_label: at most 65535 images per call");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES,
                "tss_pseudo_label: num_classes %d outside [1, %d]", num_classes, TSS_MAX_CLASSES);
    TSS_REQUIRE(mode >= TSS_PL_UP_SOFTMAX && mode <= TSS_PL_PROB_UP_SOFTMAX, TSS_E_ARG, "tss_pseudo_label: bad mode %d", mode);
    TSS_REQUIRE(flip >= TSS_FLIP_NONE && flip <= TSS_FLIP_WIDTH, TSS_E_ARG, "tss_pseudo_label: bad flip %d", flip);
    TSS_REQUIRE(flip == TSS_FLIP_NONE || src_flip != nullptr, TSS_E_ARG, "tss_pseudo_label: flip without src_flip");
    TSS_REQUIRE(ignore_index >= 0 && ignore_index <= 255, TSS_E_ARG, "tss_pseudo_label: ignore_index outside uint8");
    if (n_img == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    PLArgs a;
    a.src = src;
    a.src2 = src_flip;
    a.label = label_out;
    a.conf16 = reinterpret_cast<__half*>(conf16_out);
    a.conf32 = conf32_out;
    a.label_thr = label_thr_out;
    a.gt = nullptr;
    a.hist = nullptr;
    a.C = num_classes; a.h = h; a.w = w; a.H = H; a.W = W;
    a.sy = scale_of(h, H);
    a.sx = scale_of(w, W);
    a.mode = mode;
    a.flip = flip;
    fill_threshold(a, num_classes, thr_host, keep_host, ignore_index);
    auto al = [](const void* p, uintptr_t m) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & (m - 1)) == 0; };
    a.vec_ok = (W % 4 == 0) && al(label_out, 4) && al(conf16_out, 8) && al(conf32_out, 16) && al(label_thr_out, 4);
    a.TY = 0; a.NR = 0;
    if (hist != nullptr) {
        // the confusion matrix: fused into the register-resident kernel where that one runs (cooperative row loads,
        // per-CTA counts below 2^32), else the histogram kernel over the labels just written
        const bool fast = !g_pl_force_tile && !(h == H && w == W) && flip == TSS_FLIP_NONE &&
                          (mode == TSS_PL_UP_SOFTMAX || mode == TSS_PL_UP_ARGMAX) && (W % 2 == 0) && W >= 2 && al(label_out, 2) &&
                          al(conf16_out, 4) && al(conf32_out, 8) && al(label_thr_out, 2) && al(gt, 2) &&
                          (num_classes == 13 || num_classes == 19) && static_cast<int64_t>(num_classes) * h * w < (1ll << 30) &&
                          (a.sx * 63.0f + 3.0f) < 32.0f && !(flags & (TSS_PL_FLAG_ROWS2 | TSS_PL_FLAG_REGS96)) &&
                          n_img * static_cast<int64_t>(H) * W <= (1ll << 38);
        if (!fast) {
            TSS_REQUIRE(label_out != nullptr, TSS_E_ARG, "tss_pseudo_label_hist: this shape needs label_out for the histogram pass");
            const int rc = pseudo_label_impl(src, src_flip, n_img, num_classes, h, w, H, W, mode, flip, label_out, conf16_out,
                                             conf32_out, label_thr_out, thr_host, keep_host, ignore_index, nullptr, nullptr, flags, stream);
            if (rc != TSS_OK) return rc;
            return tss_confusion(label_out, gt, TSS_U8, n_img * static_cast<int64_t>(H) * W, num_classes, hist, nullptr, TSS_HIST_AUTO, stream);
        }
        a.gt = gt;
        a.hist = reinterpret_cast<unsigned long long*>(hist);
    }

    if (h == H && w == W && flip == TSS_FLIP_NONE && !g_pl_force_tile) {
        // full-resolution input: straight 128-bit loads, no staging
        a.vec_ok = a.vec_ok && ((static_cast<int64_t>(H) * W) % 4 == 0) && al(src, 16);
        if (num_classes <= 13) return num_classes == 13 ? launch_direct<13>(a, n_img, st) : launch_direct<13>(a, n_img, st);
        if (num_classes <= 19) return launch_direct<19>(a, n_img, st);
        return launch_direct<TSS_MAX_CLASSES>(a, n_img, st);
    }

    // register-resident fast path: logits in, no flip, even width, aligned outputs, C in {13, 19}
    if (!g_pl_force_tile && flip == TSS_FLIP_NONE && (mode == TSS_PL_UP_SOFTMAX || mode == TSS_PL_UP_ARGMAX) &&
        (W % 2 == 0) && W >= 2 && al(label_out, 2) && al(conf16_out, 4) && al(conf32_out, 8) && al(label_thr_out, 2) &&
        (num_classes == 13 || num_classes == 19) && static_cast<int64_t>(num_classes) * h * w < (1ll << 30)) {
        if (num_classes == 13)
            return mode == TSS_PL_UP_SOFTMAX ? launch_up<13, TSS_PL_UP_SOFTMAX>(a, n_img, flags, st)
                                             : launch_up<13, TSS_PL_UP_ARGMAX>(a, n_img, flags, st);
        return mode == TSS_PL_UP_SOFTMAX ? launch_up<19, TSS_PL_UP_SOFTMAX>(a, n_img, flags, st)
                                         : launch_up<19, TSS_PL_UP_ARGMAX>(a, n_img, flags, st);
    }

    // tile height: the largest of {64,32,16,8} whose staged rows fit the shared-memory budget
    const int planes = flip != TSS_FLIP_NONE ? 2 : 1;
    const size_t budget_soft = 100 * 1024, budget_hard = 220 * 1024;
    int best_ty = 0, best_nr = 0;
    size_t best_sm = 0;
    const int tys[4] = {64, 32, 16, 8};
    for (int pass = 0; pass < 2 && !best_ty; ++pass) {
        for (int k = 0; k < 4 && !best_ty; ++k) {
            const int TY = tys[k];
            int nr = 1;
            for (int y0 = 0; y0 < H; y0 += TY) {  // exact row span with the kernel's own fp32 arithmetic
                const int ylast = (y0 + TY - 1 < H ? y0 + TY - 1 : H - 1);
                int rb = static_cast<int>(a.sy * static_cast<float>(y0));
                if (rb > h - 1) rb = h - 1;
                int i0 = static_cast<int>(a.sy * static_cast<float>(ylast));
                if (i0 > h - 1) i0 = h - 1;
                const int i1 = i0 + (i0 < h - 1 ? 1 : 0);
                if (i1 - rb + 1 > nr) nr = i1 - rb + 1;
            }
            const size_t sm = static_cast<size_t>(planes) * num_classes * nr * kTX * sizeof(float);
            if (sm <= (pass == 0 ? budget_soft : budget_hard)) {
                best_ty = TY; best_nr = nr; best_sm = sm;
            }
        }
    }
    TSS_REQUIRE(best_ty != 0, TSS_E_ARG, "tss_pseudo_label: tile does not fit shared memory (C=%d, scale %f)", num_classes, a.sy);
    a.TY = best_ty;
    a.NR = best_nr;
    if (num_classes <= 13) return launch_tile<13>(a, n_img, best_sm, st);
    if (num_classes <= 19) return launch_tile<19>(a, n_img, best_sm, st);
    return launch_tile<TSS_MAX_CLASSES>(a, n_img, best_sm, st);
}
}  // namespace tss

extern "C" int tss_pseudo_label(const float* src, const float* src_flip, int64_t n_img, int num_classes, int h, int w,
                                int H, int W, int mode, int flip, uint8_t* label_out, uint16_t* conf16_out,
                                float* conf32_out, uint8_t* label_thr_out, const float* thr_host,
                                const uint8_t* keep_host, int ignore_index, int flags, void* stream) {
    return tss::pseudo_label_impl(src, src_flip, n_img, num_classes, h, w, H, W, mode, flip, label_out, conf16_out, conf32_out,
                                  label_thr_out, thr_host, keep_host, ignore_index, nullptr, nullptr, flags, stream);
}

extern "C" int tss_pseudo_label_hist(const float* src, const float* src_flip, int64_t n_img, int num_classes, int h, int w,
                                     int H, int W, int mode, int flip, uint8_t* label_out, uint16_t* conf16_out,
                                     float* conf32_out, uint8_t* label_thr_out, const float* thr_host,
                                     const uint8_t* keep_host, int ignore_index, const uint8_t* gt, int64_t* hist, int flags,
                                     void* stream) {
    using namespace tss;
    TSS_REQUIRE(gt != nullptr && hist != nullptr, TSS_E_ARG, "tss_pseudo_label_hist: null gt / hist");
    return pseudo_label_impl(src, src_flip, n_img, num_classes, h, w, H, W, mode, flip, label_out, conf16_out, conf32_out,
                             label_thr_out, thr_host, keep_host, ignore_index, gt, hist, flags, stream);
}

extern "C" int tss_softmax_nchw(const float* logits, float* probs, int64_t n_img, int num_classes, int64_t plane,
                                void* stream) {
    using namespace tss;
    TSS_REQUIRE(logits && probs && n_img >= 0 && plane > 0, TSS_E_ARG, "tss_softmax_nchw: bad arguments");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES, "tss_softmax_nchw: num_classes %d", num_classes);
    if (n_img == 0) return TSS_OK;
    const int64_t total = n_img * plane;
    int64_t blocks = (total + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    if (num_classes <= 13) softmax_nchw_kernel<13><<<static_cast<int>(blocks), 256, 0, as_stream(stream)>>>(logits, probs, n_img, num_classes, plane);
    else if (num_classes <= 19) softmax_nchw_kernel<19><<<static_cast<int>(blocks), 256, 0, as_stream(stream)>>>(logits, probs, n_img, num_classes, plane);
    else softmax_nchw_kernel<TSS_MAX_CLASSES><<<static_cast<int>(blocks), 256, 0, as_stream(stream)>>>(logits, probs, n_img, num_classes, plane);
    TSS_LAUNCH_CHECK("tss_softmax_nchw");
    return TSS_OK;
}

extern "C" int tss_upsample_bilinear(const float* src, float* dst, int64_t planes, int h, int w, int H, int W, void* stream) {
    using namespace tss;
    TSS_REQUIRE(src && dst && planes >= 0 && h > 0 && w > 0 && H > 0 && W > 0, TSS_E_ARG, "tss_upsample_bilinear: bad arguments");
    TSS_REQUIRE(H <= 65535, TSS_E_ARG, "tss_upsample_bilinear: H too large");
    if (planes == 0) return TSS_OK;
    dim3 grid((W + 255) / 256, H, static_cast<unsigned>(planes < 64 ? planes : 64));
    upsample_kernel<<<grid, 256, 0, as_stream(stream)>>>(src, dst, planes, h, w, H, W, scale_of(h, H), scale_of(w, W));
    TSS_LAUNCH_CHECK("tss_upsample_bilinear");
    return TSS_OK;
}

extern "C" int tss_threshold_labels(const uint8_t* label_in, const uint16_t* conf16, int64_t n, int num_classes,
                                    const float* thr_host, const uint8_t* keep_host, int ignore_index,
                                    uint8_t* label_out, void* stream) {
    using namespace tss;
    TSS_REQUIRE(label_in && label_out && n >= 0, TSS_E_ARG, "tss_threshold_labels: bad arguments");
    TSS_REQUIRE(thr_host == nullptr || conf16 != nullptr, TSS_E_ARG, "tss_threshold_labels: thresholds without conf");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES, "tss_threshold_labels: num_classes %d", num_classes);
    TSS_REQUIRE(ignore_index >= 0 && ignore_index <= 255, TSS_E_ARG, "tss_threshold_labels: ignore_index outside uint8");
    if (n == 0) return TSS_OK;
    PLArgs a{};
    a.C = num_classes;
    fill_threshold(a, num_classes, thr_host, keep_host, ignore_index);
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    threshold_kernel<<<static_cast<int>(blocks), 256, 0, as_stream(stream)>>>(label_in, reinterpret_cast<const __half*>(conf16), n, a, label_out);
    TSS_LAUNCH_CHECK("tss_threshold_labels");
    return TSS_OK;
}

extern "C" int tss_gather_nearest_u8(const uint8_t* src, uint8_t* dst, int64_t planes, int h, int w, int H, int W,
                                     const int32_t* ytab, const int32_t* xtab, void* stream) {
    using namespace tss;
    TSS_REQUIRE(src && dst && ytab && xtab && planes >= 0 && h > 0 && w > 0 && H > 0 && W > 0, TSS_E_ARG,
                "tss_gather_nearest_u8: bad arguments");
    TSS_REQUIRE(H <= 65535, TSS_E_ARG, "tss_gather_nearest_u8: H too large");
    if (planes == 0) return TSS_OK;
    dim3 grid((W + 255) / 256, H, static_cast<unsigned>(planes < 64 ? planes : 64));
    gather_nearest_kernel<<<grid, 256, 0, as_stream(stream)>>>(src, dst, planes, h, w, H, W, ytab, xtab);
    TSS_LAUNCH_CHECK("tss_gather_nearest_u8");
    return TSS_OK;
}
