// K3 -- prototype mask assembly: sigmoid(proto @ coef^T) -> bilinear upsample -> box crop ->
// per-pixel argmax and > 0.5, written once as bool [H, W, m] + int64 [H, W].
//
// Replaces utils_bbox.py:223 (matmul + sigmoid), :227-229 (permute / F.interpolate / permute),
// :154-168 (crop), :236 (argmax), :241 (threshold).  The reference materialises four
// [H, W, m] float tensors (121 MB each at 550^2, m = 100); here nothing but the outputs ever
// reaches HBM.
//
// Work decomposition ("strips").  The bool output of one image is one contiguous byte stream of
// H*W*m bytes (m innermost), so R full output rows are one contiguous byte range.  A work unit
// is (image, strip of R rows); CTAs pull units from an atomic counter.
//   1. zero-fill: the strip's byte range and its int64 arg-max rows are streamed out as 16-byte
//      zero stores (pure write stream -- this is where the HBM bytes go);
//   2. patch: only for detections whose crop rectangle meets the strip, the sigmoid map
//      S[det][footprint] is computed for the prototype pixels under (box  x  strip), every pixel
//      of (box  x  strip) interpolates it, and the bytes that are > 0.5 are set with byte stores
//      that land in L2 on top of the zeros written a few microseconds earlier (one DRAM write per
//      sector).  A pixel is always owned by the same thread ((row*W + col) mod 256), so the
//      running arg-max lives in shared memory without atomics and detections are visited in
//      ascending index ("first max wins", as torch.argmax).
// Work is proportional to box area; HBM traffic is the algorithmic minimum: one write of every
// output byte plus the prototype pixels that boxes touch.
// Bound: HBM write bandwidth (302 500*m + 2 420 000 bytes per image at 550^2).
#include "isb_common.cuh"

namespace isb {

constexpr int kMaskThreads = 256;
constexpr int kMaskWarps = kMaskThreads / 32;
constexpr int kStripRowsMax = 8;
constexpr int kFootH = 6;          // prototype rows a strip may touch
constexpr int kSFloats = 2048;     // sigmoid-map chunk (8 KB)
constexpr int kZeroBytes = 8192;   // zero source of the bulk (TMA) stores
constexpr int kStripPixMax = 16384;

struct MaskGeom {
  int rows;            // output rows per strip
  int strips_per_img;
  float scale_y, scale_x;
  int list_cap;        // detections per image upper bound
};

// per-column bilinear table entry: source index pair and weight (image independent)
struct ColTap { int32_t idx; float l1; };  // idx = ix0 | (ix1 - ix0) << 16

struct MaskSmem {
  size_t off_zero, off_best, off_arg, off_s, off_xtab, off_bounds, off_list, off_soff, off_pa, off_fw, total;
};
__host__ __device__ inline MaskSmem mask_smem(int rows, int out_w, int list_cap) {
  MaskSmem L;
  size_t o = 0;
  const size_t npix = (size_t)rows * out_w;
  L.off_zero = o;   o += kZeroBytes;
  L.off_best = o;   o += npix * 4;
  L.off_arg = o;    o += ((npix * 2 + 15) & ~(size_t)15);
  L.off_s = o;      o += (size_t)kSFloats * 4;
  L.off_xtab = o;   o += (size_t)out_w * sizeof(ColTap);
  o = (o + 15) & ~(size_t)15;
  L.off_bounds = o; o += (size_t)list_cap * sizeof(DetBounds);
  L.off_list = o;   o += (size_t)list_cap * 4;
  L.off_soff = o;   o += (size_t)(list_cap + 1) * 4;
  L.off_pa = o;     o += (size_t)list_cap * 2;
  L.off_fw = o;     o += (size_t)list_cap * 2;
  L.total = (o + 15) & ~(size_t)15;
  return L;
}

// Zero [g, g+nbytes): the 16-byte aligned body goes out as bulk (TMA) stores from the CTA's zero buffer, issued
// by thread 0 and completing asynchronously; the < 16-byte head and tail are plain byte stores.
__device__ __forceinline__ void cta_zero_async(uint8_t* g, int64_t nbytes, const uint8_t* zsrc, int tid) {
  int head = (int)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15);
  if (head > nbytes) head = (int)nbytes;
  const int64_t body = (nbytes - head) & ~(int64_t)15;
  const int tail = (int)(nbytes - head - body);
  if (tid < head) g[tid] = 0;
  if (tid >= 32 && tid - 32 < tail) g[head + body + (tid - 32)] = 0;
  if (tid == 0) {
    uint8_t* d = g + head;
    for (int64_t off = 0; off < body; off += kZeroBytes)
      bulk_s2g(d + off, zsrc, (uint32_t)(body - off < kZeroBytes ? body - off : kZeroBytes));
  }
}

__global__ void __launch_bounds__(kMaskThreads, 4)
mask_assemble_kernel(MaskParams p, MaskGeom g, int* __restrict__ work_counter) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ int s_unit;
  __shared__ int s_wcount[kMaskWarps];
  __shared__ int s_nlist;

  const int H = p.out_h, W = p.out_w;
  const MaskSmem L = mask_smem(g.rows, W, g.list_cap);
  const uint8_t* s_zero = smem_raw + L.off_zero;
  float* s_best = reinterpret_cast<float*>(smem_raw + L.off_best);
  uint16_t* s_arg = reinterpret_cast<uint16_t*>(smem_raw + L.off_arg);
  float* s_S = reinterpret_cast<float*>(smem_raw + L.off_s);
  ColTap* s_xtab = reinterpret_cast<ColTap*>(smem_raw + L.off_xtab);
  DetBounds* s_bounds = reinterpret_cast<DetBounds*>(smem_raw + L.off_bounds);
  int* s_list = reinterpret_cast<int*>(smem_raw + L.off_list);
  int* s_soff = reinterpret_cast<int*>(smem_raw + L.off_soff);
  int16_t* s_pa = reinterpret_cast<int16_t*>(smem_raw + L.off_pa);
  int16_t* s_fw = reinterpret_cast<int16_t*>(smem_raw + L.off_fw);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int total_units = p.batch * g.strips_per_img;
  int cached_img = -1;

  // ---- once per CTA: zero source of the bulk stores, per-column bilinear taps -------------------
  for (int i = tid; i < kZeroBytes / 16; i += kMaskThreads) reinterpret_cast<uint4*>(smem_raw + L.off_zero)[i] = make_uint4(0, 0, 0, 0);
  for (int x = tid; x < W; x += kMaskThreads) {
    int ix0, ix1;
    float l0, l1;
    bilinear_src(x, g.scale_x, p.proto_w, ix0, ix1, l0, l1);
    ColTap t;
    t.idx = ix0 | ((ix1 - ix0) << 16);
    t.l1 = l1;
    s_xtab[x] = t;
  }
  fence_proxy_async();  // the zero buffer is read by the async proxy

  while (true) {
    __syncthreads();  // everyone is done with the previous unit's shared memory (and s_unit)
    if (tid == 0) s_unit = atomicAdd(work_counter, 1);
    __syncthreads();
    const int unit = s_unit;
    if (unit >= total_units) break;
    const int b = unit / g.strips_per_img;
    const int strip = unit - b * g.strips_per_img;
    const int m = p.counts ? p.counts[b] : p.m_fixed;
    if (m <= 0) continue;  // nothing to write for this image (the reference returns None)
    const int y0 = strip * g.rows;
    const int y1 = min(y0 + g.rows, H);
    uint8_t* out_img = p.masks + (size_t)b * p.mask_slot_bytes;
    int64_t* arg_img = p.masks_arg + (size_t)b * H * W;

    // ---- 1. zero-fill (the HBM write stream), asynchronous -----------------------------------
    cta_zero_async(out_img + (int64_t)y0 * W * m, (int64_t)(y1 - y0) * W * m, s_zero, tid);
    cta_zero_async(reinterpret_cast<uint8_t*>(arg_img + (int64_t)y0 * W), (int64_t)(y1 - y0) * W * 8, s_zero, tid);
    if (tid == 0) bulk_commit();

    // ---- detection bounds of this image -> shared memory (once per image change) ----------
    if (b != cached_img) {
      const DetBounds* gb = p.bounds + (size_t)b * p.det_stride;
      for (int j = tid; j < m; j += kMaskThreads) s_bounds[j] = gb[j];
      cached_img = b;
    }
    if (tid == 0) s_nlist = 0;
    __syncthreads();  // s_bounds / s_nlist visible

    // ---- list the detections whose crop rectangle meets the strip (ordered) ----------------
    for (int base = 0; base < m; base += kMaskThreads) {
      const int j = base + tid;
      bool hit = false;
      DetBounds bd;
      if (j < m) {
        bd = s_bounds[j];
        hit = bd.ylo < y1 && bd.yhi > y0 && bd.xlo < bd.xhi;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, hit);
      if (lane == 0) s_wcount[warp] = __popc(bal);
      __syncthreads();
      int off = s_nlist, tot = 0;
#pragma unroll
      for (int w = 0; w < kMaskWarps; ++w) {
        const int c = s_wcount[w];
        if (w < warp) off += c;
        tot += c;
      }
      if (hit) {
        const int pos = off + __popc(bal & ((1u << lane) - 1u));
        const int pa = s_xtab[bd.xlo].idx & 0xffff;
        const ColTap tb = s_xtab[bd.xhi - 1];
        const int pb = (tb.idx & 0xffff) + (tb.idx >> 16);
        s_list[pos] = j;
        s_pa[pos] = (int16_t)pa;
        s_fw[pos] = (int16_t)(pb - pa + 1);
      }
      __syncthreads();
      if (tid == 0) s_nlist += tot;
      __syncthreads();
    }
    const int nlist = s_nlist;
    if (nlist == 0) continue;

    // ---- 2. patch ------------------------------------------------------------------------------
    const int nrows = y1 - y0;
    const int npix = nrows * W;
    for (int q = tid; q < npix; q += kMaskThreads) { s_best[q] = 0.0f; s_arg[q] = 0; }

    int py_min, py_max, t0, t1;
    float f0, f1;
    bilinear_src(y0, g.scale_y, p.proto_h, py_min, t1, f0, f1);
    bilinear_src(y1 - 1, g.scale_y, p.proto_h, t0, py_max, f0, f1);
    const int FH = py_max - py_min + 1;
    if (FH > kFootH) __trap();  // host strip sizing guarantees this never happens
    const float* proto_img = p.proto + (size_t)b * p.proto_h * p.proto_w * ISB_NUM_COEF;
    const float* coef_img = p.coef + (size_t)b * p.det_stride * ISB_NUM_COEF;

    bool zero_landed = false;
    int c0 = 0;
    while (c0 < nlist) {
      // chunk = as many listed detections as fit the S buffer (thread 0 lays it out)
      __syncthreads();  // previous chunk's S consumed; s_best init done
      if (tid == 0) {
        int used = 0, li = c0;
        s_soff[0] = 0;
        while (li < nlist) {
          const int need = FH * s_fw[li];
          if (used + need > kSFloats && li > c0) break;
          if (need > kSFloats) __trap();  // proto_w <= kSFloats / kFootH is enforced on the host
          used += need;
          ++li;
          s_soff[li - c0] = used;
        }
        s_nlist = li;  // reuse as "chunk end"
      }
      __syncthreads();
      const int c1 = s_nlist;
      const int nc = c1 - c0;
      const int items = s_soff[nc];
      // ---- S[det][fy][fx] = sigmoid(proto . coef) ------------------------------------------
      for (int t = tid; t < items; t += kMaskThreads) {
        int lo = 0, hi = nc - 1;  // last li with s_soff[li] <= t
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (s_soff[mid] <= t) lo = mid; else hi = mid - 1;
        }
        const int j = s_list[c0 + lo];
        const int pa = s_pa[c0 + lo], fw = s_fw[c0 + lo];
        const int f = t - s_soff[lo];
        const int fy = f / fw, fx = f - fy * fw;
        const float4* pp = reinterpret_cast<const float4*>(
            proto_img + ((size_t)(py_min + fy) * p.proto_w + (pa + fx)) * ISB_NUM_COEF);
        const float4* cf = reinterpret_cast<const float4*>(coef_img + (size_t)j * ISB_NUM_COEF);
        float acc = 0.0f;
#pragma unroll
        for (int q = 0; q < ISB_NUM_COEF / 4; ++q) {
          const float4 a = __ldg(pp + q), c = __ldg(cf + q);
          acc = fmaf(a.x, c.x, acc);
          acc = fmaf(a.y, c.y, acc);
          acc = fmaf(a.z, c.z, acc);
          acc = fmaf(a.w, c.w, acc);
        }
        s_S[t] = sigmoidf_ref(acc);
      }
      if (!zero_landed) {  // the zero stream must be complete before bytes are set on top of it
        if (tid == 0) { bulk_wait0(); fence_proxy_async_all(); }
        zero_landed = true;
      }
      __syncthreads();
      // ---- interpolate, threshold, arg-max: warp w owns row w of the strip, lanes stride the box width;
      //      a warp visits the detections in ascending index, so "first max wins" like torch.argmax -------
      for (int r = warp; r < nrows; r += kMaskWarps) {
        const int y = y0 + r;
        int iy0, iy1;
        float ly0, ly1;
        bilinear_src(y, g.scale_y, p.proto_h, iy0, iy1, ly0, ly1);
        float* brow = s_best + r * W;
        uint16_t* arow = s_arg + r * W;
        uint8_t* orow = out_img + (int64_t)y * W * m;
        for (int li = 0; li < nc; ++li) {
          const int j = s_list[c0 + li];
          const DetBounds bd = s_bounds[j];
          if (y < bd.ylo || y >= bd.yhi) continue;  // warp-uniform
          const int pa = s_pa[c0 + li], fw = s_fw[c0 + li];
          const float* S0 = s_S + s_soff[li] + (iy0 - py_min) * fw - pa;
          const float* S1 = s_S + s_soff[li] + (iy1 - py_min) * fw - pa;
          for (int x = bd.xlo + lane; x < bd.xhi; x += 32) {
            const ColTap tb = s_xtab[x];
            const int ix0 = tb.idx & 0xffff, ix1 = ix0 + (tb.idx >> 16);
            const float lx1 = tb.l1, lx0 = 1.0f - lx1;
            const float top = lx0 * S0[ix0] + lx1 * S0[ix1];
            const float bot = lx0 * S1[ix0] + lx1 * S1[ix1];
            const float v = ly0 * top + ly1 * bot;
            if (v > 0.5f) orow[(int64_t)x * m + j] = 1;
            if (v > brow[x]) { brow[x] = v; arow[x] = (uint16_t)j; }
          }
          __syncwarp();  // lanes own different pixels in the next detection
        }
      }
      c0 = c1;
    }
    __syncthreads();
    // arg-max rows: only pixels whose arg-max is not detection 0 differ from the zeros written above
    for (int q = tid; q < npix; q += kMaskThreads) {
      const int a = s_arg[q];
      if (a) arg_img[(int64_t)y0 * W + q] = (int64_t)a;
    }
  }
  if (tid == 0) bulk_wait0();  // the zero buffer must outlive every bulk store that reads it
}

// ---- host side ----------------------------------------------------------------------------
static int pick_rows(int in_size, int out_size, int rows_max, int foot_cap) {
  const float scale = (float)in_size / (float)out_size;
  int t = rows_max;
  while (t > 1 && (int)floorf(scale * (float)(t - 1)) + 3 > foot_cap) --t;
  return t;
}

static constexpr int kMaskSmemMax = 200 * 1024;

int configure_masks(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(mask_assemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaskSmemMax));
  return ISB_OK;
}

int launch_masks(isb_handle* h, const MaskParams& p, cudaStream_t s) {
  if (p.proto_w * kFootH > kSFloats || p.proto_w >= 65536)
    ISB_FAIL(h, ISB_ERR_LIMIT, "mask assembly supports prototype width <= %d, got %d", kSFloats / kFootH, p.proto_w);
  if (p.out_w > kStripPixMax) ISB_FAIL(h, ISB_ERR_LIMIT, "mask assembly supports output width <= %d, got %d", kStripPixMax, p.out_w);
  MaskGeom g;
  g.rows = pick_rows(p.proto_h, p.out_h, kStripRowsMax, kFootH);
  while (g.rows > 1 && g.rows * p.out_w > kStripPixMax) --g.rows;
  g.strips_per_img = (p.out_h + g.rows - 1) / g.rows;
  g.scale_y = (float)p.proto_h / (float)p.out_h;
  g.scale_x = (float)p.proto_w / (float)p.out_w;
  g.list_cap = p.m_fixed >= 0 ? (p.m_fixed > 0 ? p.m_fixed : 1) : p.det_stride;
  MaskSmem L = mask_smem(g.rows, p.out_w, g.list_cap);
  // prefer >= 4 resident CTAs per SM (<= 55 KB each) as long as a strip keeps >= 2 rows
  while ((L.total > (size_t)kMaskSmemMax && g.rows > 1) || (L.total > 55 * 1024 && g.rows > 2)) {
    --g.rows;
    g.strips_per_img = (p.out_h + g.rows - 1) / g.rows;
    L = mask_smem(g.rows, p.out_w, g.list_cap);
  }
  if (L.total > (size_t)kMaskSmemMax) ISB_FAIL(h, ISB_ERR_LIMIT, "mask kernel shared memory %zu exceeds %d", L.total, kMaskSmemMax);
  const int64_t total_units = (int64_t)p.batch * g.strips_per_img;
  if (total_units == 0) return ISB_OK;
  int per_sm = (int)((220 * 1024) / (L.total + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 4) per_sm = 4;
  int64_t grid = (int64_t)h->num_sms * per_sm;
  if (grid > total_units) grid = total_units;
  int* counter = h->work_counter + h->ws_lane;
  ISB_CUDA(h, cudaMemsetAsync(counter, 0, sizeof(int), s));
  mask_assemble_kernel<<<(unsigned)grid, kMaskThreads, L.total, s>>>(p, g, counter);
  ISB_LAUNCH_CHECK(h, "mask_assemble_kernel");
  return ISB_OK;
}

}  // namespace isb
