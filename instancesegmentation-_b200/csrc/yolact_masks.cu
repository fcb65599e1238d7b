// K3 -- prototype mask assembly: sigmoid(proto @ coef^T) -> bilinear upsample -> box crop ->
// per-pixel argmax and > 0.5, written once as bool [H, W, m] + int64 [H, W].
//
// Replaces utils_bbox.py:223 (matmul + sigmoid), :227-229 (permute / F.interpolate / permute),
// :154-168 (crop), :236 (argmax), :241 (threshold).  The reference materialises four
// [H, W, m] float tensors (121 MB each at 550^2, m = 100); here nothing but the outputs ever
// reaches HBM.
//
// Work decomposition.  The output of one image is a single contiguous byte stream of H*W*m
// bytes (m innermost).  A tile is tile_h x tile_w output pixels (8 x 64 for up-sampling
// configurations); a persistent grid of CTAs strides over all (image, tile) pairs.
//   * A tile first lists the detections whose crop rectangle meets it (ordered compaction, so
//     the argmax sees detections in ascending index and "first max wins" as torch.argmax).
//   * Empty list: the tile is all zeros -> straight 16-byte zero stores, no shared memory.
//   * Otherwise the tile's prototype footprint (<= 6 x 20 proto pixels x 32 ch) is staged in
//     shared memory, the sigmoid map S[det][footprint] is computed once for the listed
//     detections only, every pixel interpolates S for the detections that contain it, and the
//     tile's output bytes are composed in a shared-memory staging buffer whose 16-byte phase
//     matches the global address, then flushed with full 16-byte stores.
// Work is therefore proportional to box area, while HBM traffic is the algorithmic minimum:
// one write of every output byte, one read of the prototype rows that boxes touch.
// Bound: HBM write bandwidth (302 500 * m + 2 420 000 bytes per image at 550^2).
#include "isb_common.cuh"

namespace isb {

constexpr int kMaskThreads = 256;
constexpr int kTileHMax = 8;
constexpr int kTileWMax = 64;
constexpr int kFootH = 6;                  // prototype rows a tile may touch
constexpr int kFootW = 20;                 // prototype columns a tile may touch
constexpr int kFootMax = kFootH * kFootW;  // 120
constexpr int kProtoPitch = ISB_NUM_COEF + 1;  // 33: conflict-free scalar reads across footprint rows
constexpr int kSFloats = 6144;             // sigmoid-map chunk (24 KB)
constexpr int kStageMax = 64 * 1024;       // staging bytes (one tile row of 64 px x 1024 dets)

struct MaskSmemLayout {
  int stage_bytes;  // staging capacity incl. per-row 16-byte phase slack
  int list_cap;     // detections per image upper bound (max_det)
  size_t off_stage, off_s, off_proto, off_bounds, off_list, total;
};

__host__ __device__ inline MaskSmemLayout mask_layout(int tile_h, int tile_w, int max_det) {
  MaskSmemLayout L;
  int row_bytes = ((tile_w * max_det + 15) & ~15) + 16;
  int stage = tile_h * row_bytes;
  if (stage > kStageMax + 16 * kTileHMax) stage = ((kStageMax / row_bytes) > 0 ? (kStageMax / row_bytes) : 1) * row_bytes;
  L.stage_bytes = stage;
  L.list_cap = max_det;
  size_t o = 0;
  L.off_stage = o;  o += (size_t)stage;
  L.off_s = o;      o += (size_t)kSFloats * 4;
  L.off_proto = o;  o += (size_t)kFootMax * kProtoPitch * 4;
  o = (o + 15) & ~(size_t)15;
  L.off_bounds = o; o += (size_t)max_det * sizeof(DetBounds);
  L.off_list = o;   o += (size_t)max_det * sizeof(int);
  L.total = (o + 15) & ~(size_t)15;
  return L;
}

struct MaskGeom {
  int tile_h, tile_w, tiles_y, tiles_x;
  float scale_y, scale_x;
  int max_det;  // staging / list capacity
};

__device__ __forceinline__ void store_zero_row(uint8_t* g, int64_t nbytes, int lane) {
  // zero [g, g + nbytes) with one warp: byte head / 16-byte body / byte tail
  int head = (int)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15);
  if (head > nbytes) head = (int)nbytes;
  if (lane < head) g[lane] = 0;
  uint4* body = reinterpret_cast<uint4*>(g + head);
  int64_t nvec = (nbytes - head) >> 4;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (int64_t v = lane; v < nvec; v += 32) body[v] = z;
  int tail = (int)(nbytes - head - (nvec << 4));
  if (lane < tail) g[head + (nvec << 4) + lane] = 0;
}

__global__ void __launch_bounds__(kMaskThreads)
mask_assemble_kernel(MaskParams p, MaskGeom g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_wcount[kMaskThreads / 32];
  __shared__ int s_nlist;

  const MaskSmemLayout L = mask_layout(g.tile_h, g.tile_w, g.max_det);
  uint8_t* s_stage = smem_raw + L.off_stage;
  float* s_S = reinterpret_cast<float*>(smem_raw + L.off_s);
  float* s_proto = reinterpret_cast<float*>(smem_raw + L.off_proto);
  DetBounds* s_lb = reinterpret_cast<DetBounds*>(smem_raw + L.off_bounds);
  int* s_list = reinterpret_cast<int*>(smem_raw + L.off_list);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_per_img = g.tiles_y * g.tiles_x;
  const int64_t total_tiles = (int64_t)p.batch * tiles_per_img;
  const int H = p.out_h, W = p.out_w;

  for (int64_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    const int b = (int)(tile / tiles_per_img);
    const int trem = (int)(tile - (int64_t)b * tiles_per_img);
    const int m = p.counts ? p.counts[b] : p.m_fixed;
    if (m <= 0) continue;  // nothing to write for this image (reference returns None)
    const int ty = trem / g.tiles_x, tx = trem - ty * g.tiles_x;
    const int y0 = ty * g.tile_h, x0 = tx * g.tile_w;
    const int th = min(g.tile_h, H - y0), tw = min(g.tile_w, W - x0);
    uint8_t* out_img = p.masks + (size_t)b * p.mask_slot_bytes;
    int64_t* arg_img = p.masks_arg + (size_t)b * H * W;
    const DetBounds* gb = p.bounds + (size_t)b * p.det_stride;

    // ---- list the detections whose crop rectangle meets this tile (ordered) -------------
    __syncthreads();  // previous tile is done with shared memory
    if (tid == 0) s_nlist = 0;
    __syncthreads();
    for (int base = 0; base < m; base += kMaskThreads) {
      const int j = base + tid;
      DetBounds bd = {0, 0, 0, 0};
      bool hit = false;
      if (j < m) {
        bd = gb[j];
        hit = bd.xlo < x0 + tw && bd.xhi > x0 && bd.ylo < y0 + th && bd.yhi > y0;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, hit);
      if (lane == 0) s_wcount[warp] = __popc(bal);
      __syncthreads();
      int off = s_nlist, tot = 0;
#pragma unroll
      for (int w = 0; w < kMaskThreads / 32; ++w) {
        const int c = s_wcount[w];
        if (w < warp) off += c;
        tot += c;
      }
      if (hit) {
        const int slot = off + __popc(bal & ((1u << lane) - 1u));
        s_list[slot] = j;
        s_lb[slot] = bd;
      }
      __syncthreads();
      if (tid == 0) s_nlist += tot;
      __syncthreads();
    }
    const int nlist = s_nlist;

    if (nlist == 0) {
      // ---- all-zero tile: stream zeros --------------------------------------------------
      for (int r = warp; r < th; r += kMaskThreads / 32) {
        const int64_t pix = (int64_t)(y0 + r) * W + x0;
        store_zero_row(out_img + pix * m, (int64_t)tw * m, lane);
        for (int c = lane; c < tw; c += 32) arg_img[pix + c] = 0;
      }
      continue;
    }

    // ---- prototype footprint of the tile ------------------------------------------------
    int py_min, py_max, px_min, px_max, t0, t1;
    float f0, f1;
    bilinear_src(y0, g.scale_y, p.proto_h, py_min, t1, f0, f1);
    bilinear_src(y0 + th - 1, g.scale_y, p.proto_h, t0, py_max, f0, f1);
    bilinear_src(x0, g.scale_x, p.proto_w, px_min, t1, f0, f1);
    bilinear_src(x0 + tw - 1, g.scale_x, p.proto_w, t0, px_max, f0, f1);
    const int FH = py_max - py_min + 1, FW = px_max - px_min + 1, foot = FH * FW;
    if (FH > kFootH || FW > kFootW) __trap();  // host tile sizing guarantees this never happens
    {
      const float* proto_img = p.proto + (size_t)b * p.proto_h * p.proto_w * ISB_NUM_COEF;
      for (int t = tid; t < foot * (ISB_NUM_COEF / 4); t += kMaskThreads) {
        const int f = t >> 3, q = t & 7;
        const int fy = f / FW, fx = f - fy * FW;
        const float4 v = __ldg(reinterpret_cast<const float4*>(
                                   proto_img + ((size_t)(py_min + fy) * p.proto_w + (px_min + fx)) * ISB_NUM_COEF) + q);
        float* d = s_proto + f * kProtoPitch + q * 4;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
      }
    }

    const int dets_per_chunk = kSFloats / foot;
    const int row_pitch = ((tw * m + 15) & ~15) + 16;
    int rows_per_pass = L.stage_bytes / row_pitch;
    if (rows_per_pass > th) rows_per_pass = th;
    if (rows_per_pass < 1) __trap();
    const float* coef_img = p.coef + (size_t)b * p.det_stride * ISB_NUM_COEF;

    for (int r0 = 0; r0 < th; r0 += rows_per_pass) {
      const int pr = min(rows_per_pass, th - r0);
      const int npix = pr * tw;
      __syncthreads();  // staging free (previous pass flushed) and s_proto complete
      {
        uint4* z = reinterpret_cast<uint4*>(s_stage);
        const int nvec = (pr * row_pitch) >> 4;
        for (int v = tid; v < nvec; v += kMaskThreads) z[v] = make_uint4(0, 0, 0, 0);
      }
      // per-thread pixels of this pass (at most two: 512 pixels / 256 threads)
      int pr_row[2], pr_col[2], f00[2], f01[2], f10[2], f11[2], arg[2];
      float ly0[2], ly1[2], lx0[2], lx1[2], best[2];
      int stage_off[2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int q = tid + i * kMaskThreads;
        pr_row[i] = -1;
        if (q < npix) {
          const int r = q / tw, c = q - r * tw;
          pr_row[i] = r0 + r;
          pr_col[i] = c;
          int iy0, iy1, ix0, ix1;
          bilinear_src(y0 + r0 + r, g.scale_y, p.proto_h, iy0, iy1, ly0[i], ly1[i]);
          bilinear_src(x0 + c, g.scale_x, p.proto_w, ix0, ix1, lx0[i], lx1[i]);
          iy0 -= py_min; iy1 -= py_min; ix0 -= px_min; ix1 -= px_min;
          f00[i] = iy0 * FW + ix0; f01[i] = iy0 * FW + ix1;
          f10[i] = iy1 * FW + ix0; f11[i] = iy1 * FW + ix1;
          const int64_t gbyte = ((int64_t)(y0 + r0 + r) * W + x0) * m;
          stage_off[i] = r * row_pitch + (int)(gbyte & 15) + c * m;
          best[i] = 0.0f;
          arg[i] = 0;
        }
      }

      for (int c0 = 0; c0 < nlist; c0 += dets_per_chunk) {
        const int nc = min(dets_per_chunk, nlist - c0);
        __syncthreads();  // zeroing done / previous chunk's S consumed
        // ---- S[det][footprint] = sigmoid(proto . coef) ------------------------------------
        for (int t = tid; t < nc * foot; t += kMaskThreads) {
          const int li = t / foot, f = t - li * foot;
          const float4* cf = reinterpret_cast<const float4*>(coef_img + (size_t)s_list[c0 + li] * ISB_NUM_COEF);
          const float* pp = s_proto + f * kProtoPitch;
          float acc = 0.0f;
#pragma unroll
          for (int q = 0; q < ISB_NUM_COEF / 4; ++q) {
            const float4 cv = __ldg(cf + q);
            acc = fmaf(pp[4 * q + 0], cv.x, acc);
            acc = fmaf(pp[4 * q + 1], cv.y, acc);
            acc = fmaf(pp[4 * q + 2], cv.z, acc);
            acc = fmaf(pp[4 * q + 3], cv.w, acc);
          }
          s_S[t] = sigmoidf_ref(acc);
        }
        __syncthreads();
        // ---- interpolate, crop, threshold, argmax -----------------------------------------
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if (pr_row[i] < 0) continue;
          const int x = x0 + pr_col[i], y = y0 + pr_row[i];
          for (int li = 0; li < nc; ++li) {
            const DetBounds bd = s_lb[c0 + li];
            if (x >= bd.xlo && x < bd.xhi && y >= bd.ylo && y < bd.yhi) {
              const float* S = s_S + li * foot;
              const float top = lx0[i] * S[f00[i]] + lx1[i] * S[f01[i]];
              const float bot = lx0[i] * S[f10[i]] + lx1[i] * S[f11[i]];
              const float v = ly0[i] * top + ly1[i] * bot;
              const int j = s_list[c0 + li];
              if (v > 0.5f) s_stage[stage_off[i] + j] = 1;
              if (v > best[i]) { best[i] = v; arg[i] = j; }
            }
          }
        }
      }
      __syncthreads();
      // ---- flush: staging rows -> global with matching 16-byte phase -----------------------
      for (int r = warp; r < pr; r += kMaskThreads / 32) {
        const int64_t gbyte = ((int64_t)(y0 + r0 + r) * W + x0) * m;
        uint8_t* gp = out_img + gbyte;
        const uint8_t* sp = s_stage + r * row_pitch + (int)(gbyte & 15);
        const int nbytes = tw * m;
        int head = (16 - (int)(gbyte & 15)) & 15;
        if (head > nbytes) head = nbytes;
        if (lane < head) gp[lane] = sp[lane];
        const int nvec = (nbytes - head) >> 4;
        const uint4* sv = reinterpret_cast<const uint4*>(sp + head);
        uint4* gv = reinterpret_cast<uint4*>(gp + head);
        for (int v = lane; v < nvec; v += 32) gv[v] = sv[v];
        const int done = head + (nvec << 4);
        if (lane < nbytes - done) gp[done + lane] = sp[done + lane];
      }
#pragma unroll
      for (int i = 0; i < 2; ++i)
        if (pr_row[i] >= 0) arg_img[(int64_t)(y0 + pr_row[i]) * W + x0 + pr_col[i]] = (int64_t)arg[i];
    }
  }
}

// ---- host side ----------------------------------------------------------------------------
static int pick_tile(int in_size, int out_size, int tile_max, int foot_cap) {
  const float scale = (float)in_size / (float)out_size;
  int t = tile_max;
  while (t > 1 && (int)floorf(scale * (float)(t - 1)) + 3 > foot_cap) --t;
  return t;
}

int configure_masks(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(mask_assemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  return ISB_OK;
}

int launch_masks(isb_handle* h, const MaskParams& p, cudaStream_t s) {
  MaskGeom g;
  g.tile_h = pick_tile(p.proto_h, p.out_h, kTileHMax, kFootH);
  g.tile_w = pick_tile(p.proto_w, p.out_w, kTileWMax, kFootW);
  g.tiles_y = (p.out_h + g.tile_h - 1) / g.tile_h;
  g.tiles_x = (p.out_w + g.tile_w - 1) / g.tile_w;
  g.scale_y = (float)p.proto_h / (float)p.out_h;
  g.scale_x = (float)p.proto_w / (float)p.out_w;
  g.max_det = p.m_fixed >= 0 ? (p.m_fixed > 0 ? p.m_fixed : 1) : p.det_stride;
  const MaskSmemLayout L = mask_layout(g.tile_h, g.tile_w, g.max_det);
  if (L.total > 200 * 1024) ISB_FAIL(h, ISB_ERR_LIMIT, "mask kernel shared memory %zu exceeds 200 KiB", L.total);
  const int64_t total_tiles = (int64_t)p.batch * g.tiles_y * g.tiles_x;
  if (total_tiles == 0) return ISB_OK;
  int per_sm = (int)((220 * 1024) / (L.total + 1024));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 8) per_sm = 8;
  int64_t grid = (int64_t)h->num_sms * per_sm;
  if (grid > total_tiles) grid = total_tiles;
  mask_assemble_kernel<<<(unsigned)grid, kMaskThreads, L.total, s>>>(p, g);
  ISB_LAUNCH_CHECK(h, "mask_assemble_kernel");
  return ISB_OK;
}

}  // namespace isb
