// Wrapper pre-processing: inst_model/yolact/yolact.py:137-148 and
// inst_model/Mask_RCNN/mask_rcnn.py:140-148 --
//     resize_image(image, (W_in, H_in))            utils/utils.py:19-22   PIL Image.resize(..., Image.BICUBIC)
//     (np.array(.., float32) - mean) / std         utils/utils.py:32-36   float64 arithmetic, rounded to float32 by
//     transpose (2, 0, 1), [None], FloatTensor     yolact.py:145-148      the wrapper's .type(FloatTensor)
// Pillow's 8-bit resize (src/libImaging/Resample.c) is integer arithmetic: per axis a table of fixed-point
// coefficients (22 fractional bits, computed in double on the host exactly like precompute_coeffs /
// normalize_coeffs_8bpc), a horizontal pass into an 8-bit intermediate image, then a vertical pass; each output
// byte is clip8((2^21 + sum tap * coefficient) >> 22).  Both passes are reproduced bit for bit; the normalisation is
// a 3 x 256 table of float32((double(v) - mean) / std), also built on the host.
#include <math.h>

#include <new>
#include <vector>

#include "isb_common.cuh"

namespace isb {

constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kPrepThreads = 128;

struct AxisTable {
  int in_size, out_size, ksize;
  int32_t* d_bounds;  // [out][2]  first tap, tap count
  int32_t* d_coef;    // [ksize][out]  (transposed: consecutive outputs are adjacent)
};

struct PrepState {
  std::vector<AxisTable> tables;
  uint8_t* temp = nullptr;  // horizontal pass output [B, in_h, out_w, 3]
  size_t temp_bytes = 0;
  float* d_lut = nullptr;   // [3][256]
  double mean[3] = {0, 0, 0}, stdv[3] = {0, 0, 0};
  bool lut_valid = false;
};

// Resample.c bicubic_filter (Keys kernel, a = -0.5)
static double bicubic_filter(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}

// Resample.c precompute_coeffs + normalize_coeffs_8bpc for box = (0, in_size)
static int build_axis(isb_handle* h, int in_size, int out_size, AxisTable* t) {
  const double scale = (double)in_size / (double)out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 2.0 * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  std::vector<int32_t> bounds((size_t)out_size * 2), coef((size_t)ksize * out_size, 0);
  std::vector<double> w((size_t)ksize);
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      w[x] = bicubic_filter((x + xmin - center + 0.5) * ss);
      ww += w[x];
    }
    for (int x = 0; x < xmax; ++x) {
      if (ww != 0.0) w[x] /= ww;
      const double q = w[x] * (double)(1 << kPrecisionBits);
      coef[(size_t)x * out_size + xx] = w[x] < 0 ? (int32_t)(-0.5 + q) : (int32_t)(0.5 + q);
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
  }
  t->in_size = in_size;
  t->out_size = out_size;
  t->ksize = ksize;
  ISB_CUDA(h, cudaMalloc((void**)&t->d_bounds, bounds.size() * sizeof(int32_t)));
  ISB_CUDA(h, cudaMalloc((void**)&t->d_coef, coef.size() * sizeof(int32_t)));
  ISB_CUDA(h, cudaMemcpy(t->d_bounds, bounds.data(), bounds.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  ISB_CUDA(h, cudaMemcpy(t->d_coef, coef.data(), coef.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  return ISB_OK;
}

// tables are cached per (input size, output size): the wrappers resize every picture to the same network input
static int get_axis(isb_handle* h, PrepState* st, int in_size, int out_size, AxisTable* out) {
  for (const AxisTable& t : st->tables)
    if (t.in_size == in_size && t.out_size == out_size) { *out = t; return ISB_OK; }
  AxisTable t;
  memset(&t, 0, sizeof(t));
  int rc = build_axis(h, in_size, out_size, &t);
  if (rc != ISB_OK) return rc;
  st->tables.push_back(t);
  *out = t;
  return ISB_OK;
}

__device__ __forceinline__ uint8_t clip8(int acc) {
  const int v = acc >> kPrecisionBits;  // arithmetic shift, like Resample.c's clip8 look-up
  return (uint8_t)min(max(v, 0), 255);
}

// Horizontal pass (ImagingResampleHorizontal_8bpc): grid (column blocks, rows, images), thread = output column.
// The row bytes a block needs (the taps of its first to its last column) are staged in shared memory with coalesced
// loads; the coefficient table is transposed so that neighbouring columns read neighbouring words.
__global__ void __launch_bounds__(kPrepThreads)
prep_horizontal_kernel(const uint8_t* __restrict__ src, int in_h, int in_w, int out_w, int ksize,
                       const int32_t* __restrict__ bounds, const int32_t* __restrict__ coef, uint8_t* __restrict__ dst) {
  extern __shared__ uint8_t s_row[];
  const int xx0 = blockIdx.x * kPrepThreads, xx1 = min(xx0 + kPrepThreads, out_w) - 1;
  const int y = blockIdx.y, b = blockIdx.z;
  const int span0 = bounds[2 * xx0], span1 = bounds[2 * xx1] + bounds[2 * xx1 + 1];  // [span0, span1) input columns
  const uint8_t* row = src + ((size_t)b * in_h + y) * (size_t)in_w * 3;
  for (int i = threadIdx.x; i < (span1 - span0) * 3; i += kPrepThreads) s_row[i] = row[(size_t)span0 * 3 + i];
  __syncthreads();
  const int xx = xx0 + threadIdx.x;
  if (xx >= out_w) return;
  const int xmin = bounds[2 * xx], xmax = bounds[2 * xx + 1];
  int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
  const uint8_t* p = s_row + (xmin - span0) * 3;
  for (int x = 0; x < xmax; ++x) {
    const int k = coef[(size_t)x * out_w + xx];
    a0 += (int)p[3 * x] * k;
    a1 += (int)p[3 * x + 1] * k;
    a2 += (int)p[3 * x + 2] * k;
  }
  uint8_t* o = dst + (((size_t)b * in_h + y) * out_w + xx) * 3;
  o[0] = clip8(a0);
  o[1] = clip8(a1);
  o[2] = clip8(a2);
}

// Vertical pass (ImagingResampleVertical_8bpc) + normalisation + HWC -> CHW: grid (column blocks, rows, images),
// thread = output column; the row's taps and coefficients are block-uniform.
__global__ void __launch_bounds__(kPrepThreads)
prep_vertical_kernel(const uint8_t* __restrict__ src, int in_h, int out_h, int out_w, int ksize,
                     const int32_t* __restrict__ bounds, const int32_t* __restrict__ coef, const float* __restrict__ lut,
                     uint8_t* __restrict__ resized, float* __restrict__ input) {
  const int xx = blockIdx.x * kPrepThreads + threadIdx.x;
  const int yy = blockIdx.y, b = blockIdx.z;
  if (xx >= out_w) return;
  const int ymin = bounds[2 * yy], ymax = bounds[2 * yy + 1];
  int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
  const uint8_t* p = src + (((size_t)b * in_h + ymin) * out_w + xx) * 3;
  for (int y = 0; y < ymax; ++y) {
    const int k = coef[(size_t)y * out_h + yy];
    a0 += (int)p[0] * k;
    a1 += (int)p[1] * k;
    a2 += (int)p[2] * k;
    p += (size_t)out_w * 3;
  }
  const uint8_t v0 = clip8(a0), v1 = clip8(a1), v2 = clip8(a2);
  if (resized) {
    uint8_t* o = resized + (((size_t)b * out_h + yy) * out_w + xx) * 3;
    o[0] = v0;
    o[1] = v1;
    o[2] = v2;
  }
  if (input) {
    const size_t plane = (size_t)out_h * out_w;
    float* o = input + (size_t)b * 3 * plane + (size_t)yy * out_w + xx;
    o[0] = __ldg(lut + v0);
    o[plane] = __ldg(lut + 256 + v1);
    o[2 * plane] = __ldg(lut + 512 + v2);
  }
}

void destroy_preprocess(isb_handle* h) {
  PrepState* st = static_cast<PrepState*>(h->prep);
  if (!st) return;
  for (AxisTable& t : st->tables) {
    cudaFree(t.d_bounds);
    cudaFree(t.d_coef);
  }
  cudaFree(st->temp);
  cudaFree(st->d_lut);
  delete st;
  h->prep = nullptr;
}

int launch_preprocess(isb_handle* h, const uint8_t* images, int batch, int in_h, int in_w, int out_h, int out_w,
                      const double mean[3], const double stdv[3], uint8_t* resized, float* input, cudaStream_t s) {
  if (!h->prep) {
    h->prep = new (std::nothrow) PrepState();
    if (!h->prep) ISB_FAIL(h, ISB_ERR_CUDA, "preprocess: out of host memory");
  }
  PrepState* st = static_cast<PrepState*>(h->prep);
  int rc;
  AxisTable txv, tyv;
  const AxisTable *tx = nullptr, *ty = &tyv;
  if (in_w != out_w) {  // Resample.c skips an unchanged axis
    if ((rc = get_axis(h, st, in_w, out_w, &txv)) != ISB_OK) return rc;
    tx = &txv;
  }
  if ((rc = get_axis(h, st, in_h, out_h, &tyv)) != ISB_OK) return rc;  // unchanged height: unit coefficients, same bytes
  if (input) {
    bool same = st->lut_valid;
    for (int c = 0; c < 3 && same; ++c) same = st->mean[c] == mean[c] && st->stdv[c] == stdv[c];
    if (!same) {
      float lut[3 * 256];
      for (int c = 0; c < 3; ++c)
        for (int v = 0; v < 256; ++v) lut[c * 256 + v] = (float)(((double)v - mean[c]) / stdv[c]);
      if (!st->d_lut) ISB_CUDA(h, cudaMalloc((void**)&st->d_lut, sizeof(lut)));
      ISB_CUDA(h, cudaStreamSynchronize(s));  // an earlier call on this stream may still read the old table
      ISB_CUDA(h, cudaMemcpy(st->d_lut, lut, sizeof(lut), cudaMemcpyHostToDevice));
      for (int c = 0; c < 3; ++c) { st->mean[c] = mean[c]; st->stdv[c] = stdv[c]; }
      st->lut_valid = true;
    }
  }
  const uint8_t* vsrc = images;
  if (tx) {
    const size_t need = (size_t)batch * in_h * out_w * 3;
    if (need > st->temp_bytes) {
      ISB_CUDA(h, cudaStreamSynchronize(s));
      cudaFree(st->temp);
      st->temp = nullptr;
      st->temp_bytes = 0;
      ISB_CUDA(h, cudaMalloc((void**)&st->temp, need));
      st->temp_bytes = need;
    }
    // shared memory: the input columns one block of outputs can touch
    const double scale = (double)in_w / (double)out_w;
    const size_t smem = (size_t)((int)ceil(kPrepThreads * scale) + tx->ksize + 2) * 3;
    if (smem > 48 * 1024) ISB_FAIL(h, ISB_ERR_LIMIT, "preprocess: horizontal reduction %d -> %d is too strong for the staged row span", in_w, out_w);
    prep_horizontal_kernel<<<dim3((out_w + kPrepThreads - 1) / kPrepThreads, in_h, batch), kPrepThreads, smem, s>>>(
        images, in_h, in_w, out_w, tx->ksize, tx->d_bounds, tx->d_coef, st->temp);
    ISB_LAUNCH_CHECK(h, "prep_horizontal_kernel");
    vsrc = st->temp;
  }
  prep_vertical_kernel<<<dim3((out_w + kPrepThreads - 1) / kPrepThreads, out_h, batch), kPrepThreads, 0, s>>>(
      vsrc, in_h, out_h, out_w, ty->ksize, ty->d_bounds, ty->d_coef, st->d_lut, resized, input);
  ISB_LAUNCH_CHECK(h, "prep_vertical_kernel");
  return ISB_OK;
}

}  // namespace isb

extern "C" int isb_preprocess(isb_handle* h, const uint8_t* images, int32_t batch, int32_t in_h, int32_t in_w, int32_t out_h,
                              int32_t out_w, const double* mean, const double* std, uint8_t* resized, float* input,
                              isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (batch < 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: bad size");
  if (batch > 65535 || in_h > 65535 || out_h > 65535) ISB_FAIL(h, ISB_ERR_LIMIT, "preprocess: batch / height above 65535");
  if (batch == 0) return ISB_OK;
  if (!images || (!resized && !input)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: null pointer");
  if (input && (!mean || !std)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: mean / std required for the float output");
  if (input)
    for (int c = 0; c < 3; ++c)
      if (!(std[c] != 0.0)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: std[%d] is zero or NaN", c);
  ISB_CUDA(h, cudaSetDevice(h->device));
  return isb::launch_preprocess(h, images, batch, in_h, in_w, out_h, out_w, mean, std, resized, input, (cudaStream_t)stream);
}
