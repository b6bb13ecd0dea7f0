// On-device data augmentation fused into the batch gather: random shift (bilinear, edge-clamped) and
// horizontal flip, i.e. keras ImageDataGenerator(width_shift_range, height_shift_range,
// horizontal_flip, fill_mode='nearest') as configured at /root/reference/helpers.py:115-153, which the
// reference runs on the host with scipy in 4 worker threads.  One launch produces the augmented batch
// straight from the resident dataset: out[i, h, w, :] = bilinear(src[perm[cursor+i]], h + tx_i, w' + ty_i)
// with w' = W-1-w when flip_i is set (the flip is applied after the shift, as Keras does).
#include "common.cuh"

namespace fnet {

__global__ void __launch_bounds__(256) gather_shift_flip_kernel(int batch, int H, int W, int C,
                                                                const int32_t *__restrict__ perm,
                                                                const int32_t *__restrict__ state,
                                                                const float *__restrict__ tx, const float *__restrict__ ty,
                                                                const uint8_t *__restrict__ flip,
                                                                const float *__restrict__ src, float *__restrict__ dst) {
    pdl_prologue();
    const int cursor = state[1];
    const int64_t per_image = (int64_t)H * W * C;
    const int64_t total = (int64_t)batch * per_image;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int c = (int)(i % C);
        int64_t t = i / C;
        int w = (int)(t % W);
        t /= W;
        int h = (int)(t % H);
        int b = (int)(t / H);
        int pos = cursor + b;                      // position in the epoch: indexes the per-epoch random draws
        const float *img = src + (int64_t)__ldg(perm + pos) * per_image;
        int ws = __ldg(flip + pos) ? W - 1 - w : w;
        // scipy.ndimage.affine_transform(order=1, mode='nearest'): input coordinate = output + offset,
        // coordinates clamped to the image, linear interpolation between the two neighbours per axis
        float fh = fminf(fmaxf((float)h + __ldg(tx + pos), 0.f), (float)(H - 1));
        float fw = fminf(fmaxf((float)ws + __ldg(ty + pos), 0.f), (float)(W - 1));
        int h0 = (int)floorf(fh), w0 = (int)floorf(fw);
        int h1 = min(h0 + 1, H - 1), w1 = min(w0 + 1, W - 1);
        float ah = fh - (float)h0, aw = fw - (float)w0;
        float v00 = __ldg(img + ((int64_t)h0 * W + w0) * C + c), v01 = __ldg(img + ((int64_t)h0 * W + w1) * C + c);
        float v10 = __ldg(img + ((int64_t)h1 * W + w0) * C + c), v11 = __ldg(img + ((int64_t)h1 * W + w1) * C + c);
        float top = v00 + aw * (v01 - v00), bot = v10 + aw * (v11 - v10);
        dst[i] = top + ah * (bot - top);
    }
}

}  // namespace fnet

using namespace fnet;

extern "C" int fnet_gather_shift_flip(int32_t batch, int32_t h, int32_t w, int32_t c, const int32_t *perm,
                                      const int32_t *state, const float *tx, const float *ty, const uint8_t *flip,
                                      const float *src, float *dst, void *stream) {
    FNET_REQUIRE(batch >= 0 && h > 0 && w > 0 && c > 0, "fnet_gather_shift_flip: bad sizes");
    FNET_REQUIRE(perm && state && tx && ty && flip && src && dst, "fnet_gather_shift_flip: null pointer");
    if (batch == 0) return FNET_OK;
    int64_t total = (int64_t)batch * h * w * c;
    launch_pdl(gather_shift_flip_kernel, dim3(stream_grid(total, 256)), dim3(256), 0, as_stream(stream), batch, h, w, c, perm, state, tx, ty,
                                                                                   flip, src, dst);
    return check_launch("fnet_gather_shift_flip");
}
