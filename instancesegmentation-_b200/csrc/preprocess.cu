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
  int groups;         // ceil(longest tap run / 4)
  uint32_t* d_pack;   // [groups][3][out]: byte slices 0..2 of the coefficients of taps 4g..4g+3, one byte per tap (dp4a operand)
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
  // coefficient k = k2 * 65536 + k1 * 256 + k0 with k0, k1 unsigned bytes and k2 a signed byte (|k| < 2^23): the
  // horizontal pass forms sum(pixel * k) as three 4-way byte dot products per group of four taps
  // groups of four taps: from the longest tap run that occurs, not from ksize (1920 -> 544: ksize 17, 15 taps at most)
  int max_taps = 1;
  for (int xx = 0; xx < out_size; ++xx) max_taps = std::max(max_taps, bounds[2 * xx + 1]);
  const int groups = (max_taps + 3) / 4;
  std::vector<uint32_t> pack((size_t)groups * 3 * out_size, 0u);
  for (int xx = 0; xx < out_size; ++xx)
    for (int x = 0; x < bounds[2 * xx + 1]; ++x) {
      const int32_t k = coef[(size_t)x * out_size + xx];
      if (k >= (1 << 23) || k < -(1 << 23)) ISB_FAIL(h, ISB_ERR_LIMIT, "preprocess: coefficient out of the 24-bit range");
      const uint32_t slice[3] = {(uint32_t)k & 0xffu, ((uint32_t)k >> 8) & 0xffu, ((uint32_t)(k >> 16)) & 0xffu};
      for (int j = 0; j < 3; ++j) pack[((size_t)(x >> 2) * 3 + j) * out_size + xx] |= slice[j] << (8 * (x & 3));
    }
  t->groups = groups;
  ISB_CUDA(h, cudaMalloc((void**)&t->d_pack, pack.size() * sizeof(uint32_t)));
  ISB_CUDA(h, cudaMemcpy(t->d_pack, pack.data(), pack.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
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

// Horizontal pass (ImagingResampleHorizontal_8bpc): grid (blocks of T = 128 or 192 output columns, groups of kPrepRows
// rows, pictures), thread = output column.  Row after row, the input pixels the block's columns can touch -- the taps of
// its first to its last column -- arrive in shared memory in two steps: the raw interleaved bytes of row y + 4 go global ->
// shared with cp.async (LDGSTS, 16-byte chunks, one per thread) into a four-slot ring, and a row whose copy has landed
// is DE-INTERLEAVED into one of two plane buffers, one byte plane per channel (a thread owns 4 pixels = 12 bytes: four
// 32-bit shared loads, funnel shifts for the row's byte misalignment, six byte permutes, one 32-bit store per plane)
// while the row before it is reduced: one barrier per row.  In a plane the taps of a column are consecutive bytes, so a
// thread reads them as aligned words, realigns them with funnel shifts and multiplies four taps at a time with dp4a: the
// coefficient k is split on the host into k0 + 256 k1 + 65536 k2 (unsigned, unsigned, signed bytes), one packed word per
// slice and group of four taps, held in registers for all rows of the block; sum(pixel * k) = s0 + 256 s1 + 65536 s2
// exactly.  No per-byte extraction, and neighbouring columns start ~scale bytes apart in a plane, so a warp's loads fall
// into at most 32 consecutive words (no bank conflicts; the interleaved layout cost 100 M conflict wavefronts of 185 M,
// ncu).  The result goes to an intermediate 8-bit image whose row pitch is a multiple of 16 bytes, so that the vertical
// pass can read it with aligned 64-bit loads whatever the output width.  NG = groups of four taps (template:
// coefficients in registers); NG = 0: any tap count, coefficients re-read through L1.  The first and the last block of
// the whole batch, whose 16-byte chunks could leave the caller's buffer, stage through registers with guarded loads.
constexpr int kPrepRows = 16;
constexpr int kPrepMaxGroups = 6;  // register-resident coefficients up to 24 taps (reductions up to 5.5 : 1)

__device__ __forceinline__ uint32_t load_word_guarded(const uint8_t* p, const uint8_t* lo, const uint8_t* hi) {
  if (p >= lo && p + 4 <= hi) return __ldg(reinterpret_cast<const uint32_t*>(p));
  uint32_t w = 0;  // a word straddling the end of the caller's buffer: byte by byte
  for (int k = 0; k < 4; ++k)
    if (p + k >= lo && p + k < hi) w |= (uint32_t)p[k] << (8 * k);
  return w;
}

__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {  // a: unsigned bytes, b: signed bytes
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

template <int NG, int T>
__global__ void __launch_bounds__(T)
prep_horizontal_kernel(const uint8_t* __restrict__ src, size_t src_bytes, int in_h, int in_w, int out_w, int groups, int plane_cap,
                       const int32_t* __restrict__ bounds, const uint32_t* __restrict__ pack, uint8_t* __restrict__ dst, int tpitch) {
  extern __shared__ __align__(16) uint32_t s_planes[];  // [2 buffers][3 planes][plane_cap / 4 words]
  const int tid = threadIdx.x;
  const int xx0 = blockIdx.x * T, xx1 = min(xx0 + T, out_w) - 1, xx = xx0 + tid;
  const int xc = min(xx, out_w - 1);  // threads past the last column compute a duplicate and do not store
  const int y0 = blockIdx.y * kPrepRows, y1 = min(y0 + kPrepRows, in_h), b = blockIdx.z;
  const int span0 = bounds[2 * xx0], span1 = bounds[2 * xx1] + bounds[2 * xx1 + 1];  // input columns [span0, span1)
  const int units = (span1 - span0 + 3) >> 2;  // groups of 4 pixels
  const int pw = plane_cap >> 2;               // words per plane
  const int o = bounds[2 * xc] - span0;        // first tap of this column inside the planes
  const int wi = o >> 2, sh = (o & 3) * 8;
  uint32_t K[NG > 0 ? NG : 1][3];
  if (NG > 0) {
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
      for (int j = 0; j < 3; ++j) K[g][j] = g < groups ? pack[((size_t)g * 3 + j) * out_w + xc] : 0u;
  }
  const uint8_t* src_end = src + src_bytes;
  // everything about a row's address that does not depend on the row is worked out once: rows are in_w * 3 bytes apart
  const uint8_t* row0 = src + (((size_t)b * in_h + y0) * (size_t)in_w + span0) * 3;
  const unsigned row_step = (unsigned)in_w * 3u;
  const uint8_t* last = row0 + (size_t)(y1 - 1 - y0) * row_step;
  // block-uniform: every 32-bit load of every row of the block (aligned down, plus one word of slack) is inside the buffer
  const bool inside = row0 >= src + 3 && last + (size_t)units * 12 + 8 <= src_end;
  // this thread's first unit (4 pixels = 12 bytes) of the row to stage next; one pointer add per row instead of a fresh
  // (picture, row, column) address, and the shared-memory addresses of both buffers are fixed registers (the first
  // version spent 75 of its 180 warp instructions per row on staging, most of them address arithmetic)
  const uint8_t* up = row0 + (size_t)tid * 12;
  struct Unit { uint32_t w0, w1, w2, w3; int rsh; };
  auto fetch = [&](const uint8_t* p, Unit& r) {
    if (inside) {
      const uint32_t* q = reinterpret_cast<const uint32_t*>(p);
      r.w0 = __ldg(q); r.w1 = __ldg(q + 1); r.w2 = __ldg(q + 2); r.w3 = __ldg(q + 3);
    } else {
      r.w0 = load_word_guarded(p, src, src_end); r.w1 = load_word_guarded(p + 4, src, src_end);
      r.w2 = load_word_guarded(p + 8, src, src_end); r.w3 = load_word_guarded(p + 12, src, src_end);
    }
  };
  auto park = [&](const Unit& r, uint32_t* slot) {
    // f0 = [r0 g0 b0 r1]  f1 = [g1 b1 r2 g2]  f2 = [b2 r3 g3 b3]
    const uint32_t f0 = __funnelshift_r(r.w0, r.w1, r.rsh), f1 = __funnelshift_r(r.w1, r.w2, r.rsh),
                   f2 = __funnelshift_r(r.w2, r.w3, r.rsh);
    slot[0] = __byte_perm(__byte_perm(f0, f1, 0x0630), f2, 0x5210);
    slot[pw] = __byte_perm(__byte_perm(f0, f1, 0x0741), f2, 0x6210);
    slot[2 * pw] = __byte_perm(__byte_perm(f0, f1, 0x0052), f2, 0x7410);
  };
  const bool has_unit = tid < units, more_units = units > T;  // the second is block-uniform and false up to 4 : 1
  // A row is staged in two halves around the reduction of the row before it: the loads are issued first (stage_fetch),
  // their permutes and shared-memory stores come after the reduction (stage_park), so the global-memory latency is
  // covered by the thread's own ~90 instructions instead of by other warps only.
  Unit pend;
  pend.w0 = pend.w1 = pend.w2 = pend.w3 = 0u; pend.rsh = 0;
  auto stage_fetch = [&]() {
    pend.rsh = (int)(reinterpret_cast<uintptr_t>(up) & 3) * 8;  // T * 12 bytes per step: the same for every unit of the row
    if (has_unit) fetch(up - (pend.rsh >> 3), pend);
  };
  auto stage_park = [&](uint32_t* slot) {  // slot = this thread's word of the buffer's first plane
    if (has_unit) park(pend, slot);
    if (more_units) {
      const uint8_t* p = up - (pend.rsh >> 3);
#pragma unroll 1
      for (int u = tid + T; u < units; u += T) {
        Unit r;
        r.rsh = pend.rsh;
        fetch(p + (size_t)(u - tid) * 12, r);
        park(r, slot + (u - tid));
      }
    }
    up += row_step;
  };
  uint8_t* orow = dst + ((size_t)b * in_h + y0) * (size_t)tpitch + (size_t)xc * 3;
  const bool store = xx < out_w;
  auto reduce_row = [&](const uint32_t* buf) {  // buf = first word of this column's taps in the first plane
    int res[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const uint32_t* pl = buf + c * pw;
      int s0 = 1 << (kPrecisionBits - 1), s1 = 0, s2 = 0;  // the rounding term rides in the low slice
      uint32_t prev = pl[0];
      if (NG > 0) {
#pragma unroll
        for (int g = 0; g < NG; ++g) {
          const uint32_t next = pl[g + 1];
          const uint32_t f = __funnelshift_r(prev, next, sh);
          s0 = dp4a_uu(f, K[g][0], s0);
          s1 = dp4a_uu(f, K[g][1], s1);
          s2 = dp4a_us(f, K[g][2], s2);
          prev = next;
        }
      } else {
        for (int g = 0; g < groups; ++g) {
          const uint32_t next = pl[g + 1];
          const uint32_t f = __funnelshift_r(prev, next, sh);
          const uint32_t* kp = pack + (size_t)g * 3 * out_w + xc;
          s0 = dp4a_uu(f, __ldg(kp), s0);
          s1 = dp4a_uu(f, __ldg(kp + out_w), s1);
          s2 = dp4a_us(f, __ldg(kp + 2 * out_w), s2);
          prev = next;
        }
      }
      // modulo 2^32 like Resample.c's int accumulator: the slices may wrap, the total does not
      res[c] = (int)((uint32_t)s0 + ((uint32_t)s1 << 8) + ((uint32_t)s2 << 16));
    }
    if (store) {
      orow[0] = clip8(res[0]);
      orow[1] = clip8(res[1]);
      orow[2] = clip8(res[2]);
    }
    orow += tpitch;
  };
  // two rows per trip so that the buffer a row is staged into / reduced from is known at compile time
  uint32_t* const slot0 = s_planes + tid;
  uint32_t* const slot1 = slot0 + 3 * pw;
  const uint32_t* const taps0 = s_planes + wi;
  const uint32_t* const taps1 = taps0 + 3 * pw;
  // ---- asynchronous staging (every block whose 16-byte chunks lie inside the caller's buffer) --------------------------
  // ncu on the register-staged loop: the first dp4a of every row waited ~35 % of the kernel on a long scoreboard although
  // its operands come from shared memory -- the row's global loads and its tap loads share scoreboard counters, so
  // issuing the loads early did not hide their latency.  LDGSTS has no register scoreboard: the raw bytes of row y + 4 go
  // from global to a 4-slot shared-memory ring in 16-byte chunks (one per thread), completion is a cp.async group wait
  // three rows later, and the de-interleaving into planes reads the ring.
  const int raw_cap = (plane_cap * 3 + 48 + 15) & ~15;  // bytes per ring slot (launch_preprocess sizes the same)
  uint8_t* const ring = reinterpret_cast<uint8_t*>(s_planes) + ((6 * plane_cap + 15) & ~15);
  const uint8_t* const first16 = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(row0) & ~(uintptr_t)15);
  const bool inside16 = inside && first16 >= src &&
                        reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(last) & ~(uintptr_t)15) + raw_cap <= src_end;
  if (inside16) {
    // running per-thread state instead of per-row address arithmetic: the byte address of this thread's 16-byte chunk
    // before alignment (its low four bits are the row's misalignment d), the ring offsets of the slot written / read next
    const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring) + 16u * tid;
    const uint32_t ring_bytes = 4u * (uint32_t)raw_cap;
    uintptr_t g = reinterpret_cast<uintptr_t>(row0) + 16u * tid;
    uint32_t ioff = 0, poff = 0;
    int to_issue = y1 - y0;  // rows not yet issued
    const int ubytes = units * 12 + 4 + 15;
    const bool more_chunks = ((15 + ubytes) >> 4) > T;  // block-uniform, false up to ~5 : 1
    auto issue = [&]() {  // one commit group per call, empty past the last row
      if (to_issue > 0) {
        const int chunks = (int)(((uint32_t)g & 15u) + ubytes) >> 4;
        if (tid < chunks) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_s + ioff), "l"(g & ~(uintptr_t)15) : "memory");
        if (more_chunks) {
#pragma unroll 1
          for (int c = tid + T; c < chunks; c += T)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_s + ioff + 16u * (c - tid)),
                         "l"((g & ~(uintptr_t)15) + 16u * (unsigned)(c - tid)) : "memory");
        }
        g += row_step;
        ioff += (uint32_t)raw_cap;
        if (ioff == ring_bytes) ioff = 0;
      }
      --to_issue;
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    uint32_t pd = (uint32_t)(reinterpret_cast<uintptr_t>(row0) & 15);  // misalignment of the row parked next
    const uint32_t u12 = 12u * tid;
    auto park_raw = [&](uint32_t* slot) {
      Unit r;
      r.rsh = (int)(pd & 3u) * 8;
      const uint8_t* raw = ring + poff;
      if (has_unit) {
        const uint32_t* q = reinterpret_cast<const uint32_t*>(raw + ((pd + u12) & ~3u));
        r.w0 = q[0]; r.w1 = q[1]; r.w2 = q[2]; r.w3 = q[3];
        park(r, slot);
      }
      if (more_units) {
#pragma unroll 1
        for (int u = tid + T; u < units; u += T) {
          const uint32_t* q = reinterpret_cast<const uint32_t*>(raw + ((pd + 12u * u) & ~3u));
          r.w0 = q[0]; r.w1 = q[1]; r.w2 = q[2]; r.w3 = q[3];
          park(r, slot + (u - tid));
        }
      }
      pd = (pd + row_step) & 15u;
      poff += (uint32_t)raw_cap;
      if (poff == ring_bytes) poff = 0;
    };
    issue(); issue(); issue(); issue();
    asm volatile("cp.async.wait_group 3;" ::: "memory");
    __syncthreads();
    park_raw(slot0);  // row y0
    asm volatile("cp.async.wait_group 2;" ::: "memory");
    for (int y = y0; y < y1; y += 2) {
      __syncthreads();  // row y in planes 0, raw row y + 1 landed; ring slot of raw row y is free
      issue();          // raw row y + 4
      if (y + 1 < y1) park_raw(slot1);
      reduce_row(taps0);
      if (y + 1 >= y1) break;
      asm volatile("cp.async.wait_group 2;" ::: "memory");
      __syncthreads();
      issue();
      if (y + 2 < y1) park_raw(slot0);
      reduce_row(taps1);
      asm volatile("cp.async.wait_group 2;" ::: "memory");
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    return;
  }
  // ---- first / last block of the whole batch: register staging with guarded loads ----------------------------------
  stage_fetch();
  stage_park(slot0);
  for (int y = y0; y < y1; y += 2) {
    __syncthreads();  // row y staged; the other buffer is free again
    if (y + 1 < y1) stage_fetch();
    reduce_row(taps0);
    if (y + 1 >= y1) break;
    stage_park(slot1);
    __syncthreads();
    if (y + 2 < y1) stage_fetch();
    reduce_row(taps1);
    if (y + 2 < y1) stage_park(slot0);
  }
}

// Vertical pass (ImagingResampleVertical_8bpc) + normalisation + HWC -> CHW: grid (column blocks, groups of kPrepVRows
// output rows, pictures), 256 threads.  The pass is independent per byte of a row, so a thread reduces EIGHT adjacent
// bytes: one aligned 64-bit load per tap row (block-uniform taps and coefficient), eight byte extractions (PRMT) and
// eight multiply-adds -- per byte 2.6 instructions per tap instead of 4 with four bytes per thread (ncu, round 2: the
// pass ran at 76 % of the issue slots with 215 warp instructions per 128 bytes of a row).  A block covers `ppb` byte
// octets (a multiple of 3, so that its byte range starts and ends on a pixel boundary; 207 for a 550-pixel row: one
// block, 6.5 of its 7 reducing warps busy).  The clipped bytes go to shared memory; then a thread per pixel picks its
// three channels, looks the float values up and writes one coalesced run per channel plane (plus the optional 8-bit
// picture).
constexpr int kPrepVThreads = 256;
constexpr int kPrepVRows = 8;
constexpr int kPrepVMaxOctets = 222;  // <= 7 warps reduce; 222 * 8 / 3 = 592 pixels
constexpr int kPrepVMaxTaps = 32;

__global__ void __launch_bounds__(kPrepVThreads)
prep_vertical_kernel(const uint8_t* __restrict__ src, int tpitch, int in_h, int out_h, int out_w, int ppb, int ksize,
                     const int32_t* __restrict__ bounds, const int32_t* __restrict__ coef, const float* __restrict__ lut,
                     uint8_t* __restrict__ resized, float* __restrict__ input) {
  __shared__ uint2 s_out[2][kPrepVMaxOctets];
  __shared__ int32_t s_coef[kPrepVRows][kPrepVMaxTaps];  // the block's coefficients: one LDS with an immediate offset per tap
  const int tid = threadIdx.x, b = blockIdx.z;
  const int npx = (ppb * 8) / 3;        // pixels per block
  const int px0 = blockIdx.x * npx;     // first pixel of the block
  const int octet = blockIdx.x * ppb + tid;
  const bool reducer = tid < ppb && octet * 8 < tpitch;  // tpitch is a multiple of 16
  const int yy0 = blockIdx.y * kPrepVRows, yy1 = min(yy0 + kPrepVRows, out_h);
  const size_t plane = (size_t)out_h * out_w;
  const uint2* img = reinterpret_cast<const uint2*>(src + (size_t)b * in_h * (size_t)tpitch) + octet;
  const unsigned pitch8 = (unsigned)(tpitch >> 3);
  const bool staged = ksize <= kPrepVMaxTaps;  // block-uniform; longer tap runs (reductions beyond 7.5 : 1) read the table
  if (staged) {
    for (int i = tid; i < (yy1 - yy0) * ksize; i += kPrepVThreads) {
      const int r = i / ksize, y = i - r * ksize;
      s_coef[r][y] = __ldg(coef + (size_t)y * out_h + yy0 + r);
    }
    __syncthreads();
  }
  for (int yy = yy0; yy < yy1; ++yy) {
    uint2* so = s_out[(yy - yy0) & 1];
    if (reducer) {
      const int ymin = __ldg(bounds + 2 * yy), ymax = __ldg(bounds + 2 * yy + 1);
      int a[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = 1 << (kPrecisionBits - 1);
      const uint2* p = img + (size_t)(unsigned)ymin * pitch8;
      auto tap = [&](int k) {
        const uint2 w = __ldg(p);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          a[j] += (int)__byte_perm(w.x, 0u, 0x4440u + j) * k;
          a[4 + j] += (int)__byte_perm(w.y, 0u, 0x4440u + j) * k;
        }
        p += pitch8;
      };
      if (staged) {
        const int32_t* ks = s_coef[yy - yy0];
#pragma unroll 4
        for (int y = 0; y < ymax; ++y) tap(ks[y]);
      } else {
        const int32_t* kp = coef + yy;
#pragma unroll 2
        for (int y = 0; y < ymax; ++y, kp += out_h) tap(__ldg(kp));
      }
      uint2 o;
      o.x = (uint32_t)clip8(a[0]) | ((uint32_t)clip8(a[1]) << 8) | ((uint32_t)clip8(a[2]) << 16) | ((uint32_t)clip8(a[3]) << 24);
      o.y = (uint32_t)clip8(a[4]) | ((uint32_t)clip8(a[5]) << 8) | ((uint32_t)clip8(a[6]) << 16) | ((uint32_t)clip8(a[7]) << 24);
      so[tid] = o;
    }
    __syncthreads();  // one barrier per row: the buffer written two rows ago was read before the previous barrier
    for (int px = tid; px < npx; px += kPrepVThreads) {
      const int xx = px0 + px;
      if (xx >= out_w) break;
      const uint8_t* sb = reinterpret_cast<const uint8_t*>(so) + px * 3;
      const uint8_t v0 = sb[0], v1 = sb[1], v2 = sb[2];
      if (resized) {
        uint8_t* r = resized + (((size_t)b * out_h + yy) * out_w + xx) * 3;
        r[0] = v0;
        r[1] = v1;
        r[2] = v2;
      }
      if (input) {
        float* o = input + (size_t)b * 3 * plane + (size_t)yy * out_w + xx;
        o[0] = __ldg(lut + v0);
        o[plane] = __ldg(lut + 256 + v1);
        o[2 * plane] = __ldg(lut + 512 + v2);
      }
    }
  }
}

void destroy_preprocess(isb_handle* h) {
  PrepState* st = static_cast<PrepState*>(h->prep);
  if (!st) return;
  for (AxisTable& t : st->tables) {
    cudaFree(t.d_bounds);
    cudaFree(t.d_coef);
    cudaFree(t.d_pack);
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
  AxisTable tx, ty;
  // Resample.c skips an axis whose size does not change; the unit coefficients computed for it give the same bytes,
  // so both passes always run and the vertical pass always reads the aligned intermediate image.
  if ((rc = get_axis(h, st, in_w, out_w, &tx)) != ISB_OK) return rc;
  if ((rc = get_axis(h, st, in_h, out_h, &ty)) != ISB_OK) return rc;
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
  const int tpitch = (out_w * 3 + 15) & ~15;
  const size_t need = (size_t)batch * in_h * tpitch;
  if (need > st->temp_bytes) {
    ISB_CUDA(h, cudaStreamSynchronize(s));
    cudaFree(st->temp);
    st->temp = nullptr;
    st->temp_bytes = 0;
    ISB_CUDA(h, cudaMalloc((void**)&st->temp, need));
    ISB_CUDA(h, cudaMemset(st->temp, 0, need));  // the pitch padding is read (and discarded) by the vertical pass
    st->temp_bytes = need;
  }
  // shared memory of the horizontal pass: two buffers of three byte planes for the input pixels one block of output
  // columns can touch, plus the words a column reads past its last tap
  const double scale = (double)in_w / (double)out_w;
  const int groups = tx.groups, ng = groups <= kPrepMaxGroups ? groups : 0;
  // columns per block = threads per block: 128 or 192, whichever leaves fewer idle threads in the last block of a row
  // (550 columns: 5 x 128 = 640 slots against 3 x 192 = 576)
  const int waste128 = (out_w + 127) / 128 * 128 - out_w, waste192 = (out_w + 191) / 192 * 192 - out_w;
  const int ht = (waste192 < waste128 && !getenv("ISB_PREP_T128")) ? 192 : 128;
  const int plane_cap = (((int)ceil(ht * scale) + tx.ksize + 2 + 4 * (ng > 0 ? ng : groups) + 8 + 3) & ~3);
  // two plane buffers + the 4-slot ring of raw rows (same formula as in the kernel)
  const size_t smem = (((size_t)6 * plane_cap + 15) & ~(size_t)15) + 4 * (size_t)((plane_cap * 3 + 48 + 15) & ~15);
  if (smem > 96 * 1024) ISB_FAIL(h, ISB_ERR_LIMIT, "preprocess: horizontal reduction %d -> %d is too strong for the staged row span", in_w, out_w);
  const dim3 hgrid((out_w + ht - 1) / ht, (in_h + kPrepRows - 1) / kPrepRows, batch);
  const size_t src_bytes = (size_t)batch * in_h * in_w * 3;
#define ISB_PREP_HT(NGV, TV)                                                                                                       \
  do {                                                                                                                             \
    if (smem > 48 * 1024)                                                                                                          \
      ISB_CUDA(h, cudaFuncSetAttribute(prep_horizontal_kernel<NGV, TV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)); \
    prep_horizontal_kernel<NGV, TV><<<hgrid, TV, smem, s>>>(images, src_bytes, in_h, in_w, out_w, groups, plane_cap, tx.d_bounds,  \
                                                            tx.d_pack, st->temp, tpitch);                                          \
  } while (0)
#define ISB_PREP_H(NGV)          \
  do {                           \
    if (ht == 192)               \
      ISB_PREP_HT(NGV, 192);     \
    else                         \
      ISB_PREP_HT(NGV, 128);     \
  } while (0)
  switch (ng) {
    case 1: ISB_PREP_H(1); break;
    case 2: ISB_PREP_H(2); break;
    case 3: ISB_PREP_H(3); break;
    case 4: ISB_PREP_H(4); break;
    case 5: ISB_PREP_H(5); break;
    case 6: ISB_PREP_H(6); break;
    default: ISB_PREP_H(0); break;
  }
#undef ISB_PREP_H
#undef ISB_PREP_HT
  ISB_LAUNCH_CHECK(h, "prep_horizontal_kernel");
  // byte octets per block: a multiple of 3 (pixel boundaries), as even a split of the row as 222 per block allows
  const int octets = tpitch >> 3, vblocks = (octets + kPrepVMaxOctets - 1) / kPrepVMaxOctets;
  const int ppb = ((octets + vblocks - 1) / vblocks + 2) / 3 * 3;
  prep_vertical_kernel<<<dim3((octets + ppb - 1) / ppb, (out_h + kPrepVRows - 1) / kPrepVRows, batch), kPrepVThreads, 0, s>>>(
      st->temp, tpitch, in_h, out_h, out_w, ppb, ty.ksize, ty.d_bounds, ty.d_coef, st->d_lut, resized, input);
  ISB_LAUNCH_CHECK(h, "prep_vertical_kernel");
  return ISB_OK;
}

}  // namespace isb

extern "C" int isb_preprocess(isb_handle* h, const uint8_t* images, int32_t batch, int32_t in_h, int32_t in_w, int32_t out_h,
                              int32_t out_w, const double* mean, const double* std, uint8_t* resized, float* input,
                              isb_stream stream) {
  if (!h) return ISB_ERR_BAD_ARG;
  if (batch < 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: bad size");
  if (batch > 65535 || in_h > 65535 * 16 || out_h > 65535 * 8) ISB_FAIL(h, ISB_ERR_LIMIT, "preprocess: batch / height too large");
  if (batch == 0) return ISB_OK;
  if (!images || (!resized && !input)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: null pointer");
  if (input && (!mean || !std)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: mean / std required for the float output");
  if (input)
    for (int c = 0; c < 3; ++c)
      if (!(std[c] != 0.0)) ISB_FAIL(h, ISB_ERR_BAD_ARG, "preprocess: std[%d] is zero or NaN", c);
  ISB_SET_DEVICE(h);
  return isb::launch_preprocess(h, images, batch, in_h, in_w, out_h, out_w, mean, std, resized, input, (cudaStream_t)stream);
}
