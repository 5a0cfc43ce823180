// A9 full2weak (self_training.py:145-158) and process_label (cal_prototype.py:146-151,
// self_training.py:204-209).
//
// full2weak: the reference resizes every image's feature map to (int(h/4), int(w/4)) of its RandomSized
// dims (bilinear, align_corners=True), crops the RandomCrop window / 4, mirrors W when the weak view was
// flipped, and concatenates.  That materialises a resized [K, rh, rw] tensor per image and copies it
// twice more.  Here one gather kernel samples the source map at the transformed coordinates and writes
// the [N, K, TH, TW] result once; the tap arithmetic is the same fp32 sequence as ATen's
// upsample_bilinear2d (scale = (in-1)/(out-1); src = scale*dst; lambda1 = src - floor).
#include "common.cuh"

namespace tss {

struct WeakParam {  // one row of the DEVICE int32 table [N, 6]
    int rh, rw;     // resized extents int(h/4), int(w/4)
    int y0, x0;     // crop origin in the resized map
    int flip;       // mirror along W after the crop
    int pad;
};

__device__ __forceinline__ void tap_of(float scale, int dst, int n_in, int& i0, int& i1, float& l0, float& l1) {
    const float s = __fmul_rn(scale, static_cast<float>(dst));  // rounded product: never contracted into the subtraction
    i0 = min(static_cast<int>(s), n_in - 1);
    i1 = i0 + (i0 < n_in - 1 ? 1 : 0);
    l1 = __fsub_rn(s, static_cast<float>(i0));
    l0 = __fsub_rn(1.0f, l1);
}

__global__ void __launch_bounds__(256)
full2weak_kernel(const float* __restrict__ feat, int64_t K, int h, int w, const WeakParam* __restrict__ params,
                 int TH, int TW, float* __restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y % TH;
    const int64_t n = blockIdx.y / TH;
    if (x >= TW) return;
    const WeakParam p = params[n];
    const float sy = p.rh > 1 ? static_cast<float>(h - 1) / static_cast<float>(p.rh - 1) : 0.f;
    const float sx = p.rw > 1 ? static_cast<float>(w - 1) / static_cast<float>(p.rw - 1) : 0.f;
    const int xr = p.x0 + (p.flip ? (TW - 1 - x) : x);
    const int yr = p.y0 + y;
    int x0, x1, y0, y1;
    float lx0, lx1, ly0, ly1;
    tap_of(sx, xr, w, x0, x1, lx0, lx1);
    tap_of(sy, yr, h, y0, y1, ly0, ly1);
    const int64_t plane = static_cast<int64_t>(h) * w;
    const float* src = feat + n * K * plane;
    float* dst = out + ((n * K) * TH + y) * static_cast<int64_t>(TW) + x;
    const int64_t ostride = static_cast<int64_t>(TH) * TW;
    for (int64_t k = blockIdx.z; k < K; k += gridDim.z) {
        const float* q = src + k * plane;
        const float top = fmaf(lx1, __ldg(q + y0 * w + x1), lx0 * __ldg(q + y0 * w + x0));
        const float bot = fmaf(lx1, __ldg(q + y1 * w + x1), lx0 * __ldg(q + y1 * w + x0));
        dst[k * ostride] = fmaf(ly1, bot, ly0 * top);
    }
}

__global__ void __launch_bounds__(256)
one_hot_kernel(const float* __restrict__ label, int64_t P, int C, int overflow, float* __restrict__ out) {
    const int64_t n = blockIdx.y;
    for (int64_t p = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; p < P;
         p += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const float lv = label[n * P + p];
        const int id = (lv < static_cast<float>(C)) ? static_cast<int>(lv) : overflow;
        float* o = out + n * (C + 1) * P + p;
        for (int c = 0; c <= C; ++c) o[c * P] = (c == id) ? 1.f : 0.f;
    }
}

// adjoint of the align_corners=True bilinear upsample: grad_in[planes,h,w] += taps^T grad_out[planes,H,W].
// One thread per output pixel and plane group; the four weighted contributions go out as fp32 atomics
// (RED.ADD.F32), which land in L2 -- grad_in is 64x smaller than grad_out for the 1/8-stride maps.
__global__ void __launch_bounds__(256)
upsample_adjoint_kernel(const float* __restrict__ gout, float* __restrict__ gin, int64_t planes, int h, int w, int H,
                        int W, float sy, float sx) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= W) return;
    int x0, x1, y0, y1;
    float lx0, lx1, ly0, ly1;
    tap_of(sx, x, w, x0, x1, lx0, lx1);
    tap_of(sy, y, h, y0, y1, ly0, ly1);
    for (int64_t pl = blockIdx.z; pl < planes; pl += gridDim.z) {
        const float g = gout[(pl * H + y) * static_cast<int64_t>(W) + x];
        float* q = gin + pl * static_cast<int64_t>(h) * w;
        atomicAdd(q + y0 * w + x0, ly0 * lx0 * g);
        atomicAdd(q + y0 * w + x1, ly0 * lx1 * g);
        atomicAdd(q + y1 * w + x0, ly1 * lx0 * g);
        atomicAdd(q + y1 * w + x1, ly1 * lx1 * g);
    }
}

}  // namespace tss

extern "C" int tss_upsample_bilinear_adjoint(const float* grad_out, float* grad_in, int64_t planes, int h, int w,
                                             int H, int W, void* stream) {
    using namespace tss;
    TSS_REQUIRE(grad_out && grad_in && planes >= 0 && h > 0 && w > 0 && H > 0 && W > 0, TSS_E_ARG,
                "tss_upsample_bilinear_adjoint: bad arguments");
    TSS_REQUIRE(H <= 65535, TSS_E_ARG, "tss_upsample_bilinear_adjoint: H too large");
    if (planes == 0) return TSS_OK;
    cudaStream_t st = as_stream(stream);
    cudaError_t e = cudaMemsetAsync(grad_in, 0, sizeof(float) * planes * h * w, st);
    if (e != cudaSuccess) return cuda_status(e, "tss_upsample_bilinear_adjoint: memset");
    const float sy = H > 1 ? static_cast<float>(h - 1) / static_cast<float>(H - 1) : 0.f;
    const float sx = W > 1 ? static_cast<float>(w - 1) / static_cast<float>(W - 1) : 0.f;
    dim3 grid((W + 255) / 256, H, static_cast<unsigned>(planes < 64 ? planes : 64));
    upsample_adjoint_kernel<<<grid, 256, 0, st>>>(grad_out, grad_in, planes, h, w, H, W, sy, sx);
    TSS_LAUNCH_CHECK("tss_upsample_bilinear_adjoint");
    return TSS_OK;
}

extern "C" int tss_full2weak(const float* feat, int64_t n_img, int64_t channels, int h, int w,
                             const int32_t* params, int TH, int TW, float* out, void* stream) {
    using namespace tss;
    TSS_REQUIRE(feat && params && out, TSS_E_ARG, "tss_full2weak: null pointer");
    TSS_REQUIRE(n_img >= 0 && channels > 0 && h > 0 && w > 0 && TH > 0 && TW > 0, TSS_E_ARG, "tss_full2weak: bad extents");
    TSS_REQUIRE(n_img * TH <= 65535, TSS_E_ARG, "tss_full2weak: N*TH too large");
    if (n_img == 0) return TSS_OK;
    dim3 grid((TW + 255) / 256, static_cast<unsigned>(n_img * TH), static_cast<unsigned>(channels < 32 ? channels : 32));
    dim3 block(TW >= 256 ? 256 : ((TW + 31) / 32) * 32);
    grid.x = (TW + block.x - 1) / block.x;
    full2weak_kernel<<<grid, block, 0, as_stream(stream)>>>(feat, channels, h, w, reinterpret_cast<const WeakParam*>(params), TH, TW, out);
    TSS_LAUNCH_CHECK("tss_full2weak");
    return TSS_OK;
}

extern "C" int tss_one_hot_overflow(const float* label, int64_t n_img, int64_t plane, int num_classes,
                                    int overflow_channel, float* out, void* stream) {
    using namespace tss;
    TSS_REQUIRE(label && out, TSS_E_ARG, "tss_one_hot_overflow: null pointer");
    TSS_REQUIRE(n_img >= 0 && n_img <= 65535 && plane > 0, TSS_E_ARG, "tss_one_hot_overflow: bad extents");
    TSS_REQUIRE(num_classes >= 1 && num_classes <= TSS_MAX_CLASSES, TSS_E_CLASSES, "tss_one_hot_overflow: num_classes %d", num_classes);
    TSS_REQUIRE(overflow_channel >= 0 && overflow_channel <= num_classes, TSS_E_ARG,
                "tss_one_hot_overflow: overflow channel %d outside [0, %d] (the reference's scatter_ raises)", overflow_channel, num_classes);
    if (n_img == 0) return TSS_OK;
    int64_t bx = (plane + 255) / 256;
    if (bx > kNumSMs * 8) bx = kNumSMs * 8;
    dim3 grid(static_cast<unsigned>(bx), static_cast<unsigned>(n_img));
    one_hot_kernel<<<grid, 256, 0, as_stream(stream)>>>(label, plane, num_classes, overflow_channel, out);
    TSS_LAUNCH_CHECK("tss_one_hot_overflow");
    return TSS_OK;
}
