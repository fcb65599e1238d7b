// K3 -- prototype mask assembly: sigmoid(proto @ coef^T) -> bilinear upsample -> box crop ->
// per-pixel argmax and > 0.5, written once as bool [H, W, m] + int64 [H, W].
//
// Replaces utils_bbox.py:223 (matmul + sigmoid), :227-229 (permute / F.interpolate / permute),
// :154-168 (crop), :236 (argmax), :241 (threshold).  The reference materialises four
// [H, W, m] float tensors (121 MB each at 550^2, m = 100); here nothing but the outputs ever
// reaches HBM.
//
// Two kernels:
//   K3a sigmoid_map_kernel  S_j = sigmoid(proto . coef_j) for every detection j over the prototype
//       pixels under detection j's crop box only (its "footprint", plus the bilinear halo), written to
//       an L2-resident scratch (about 0.1 MB per image at the benchmark's box sizes).  Everything outside
//       a crop box is zero after :230, so nothing else is ever needed.
//   K3b mask_compose_kernel  the output of one image is ONE contiguous byte stream of H*W pixels x m
//       bytes (m innermost).  A work unit is a contiguous range of kUnitPx pixels; CTAs pull units from an
//       atomic counter and cut them into tiles of <= 512 pixels.  One thread owns one pixel of the tile:
//       it walks the detections whose box covers it in ascending index (warp-level culling first),
//       interpolates S with aten's align_corners=False arithmetic, keeps the running arg-max in
//       registers ("first max wins", as torch.argmax), and sets the pixel's > 0.5 bytes in a shared-memory
//       image of the tile (byte stride m: conflict-free for m = 100).  The finished tile leaves the SM
//       as ONE bulk (TMA) store, double-buffered, so every output byte is written exactly once, in full
//       lines, asynchronously; the int64 arg-max goes out as coalesced 8-byte stores.  Tiles stay
//       zero except for the bytes their owner warp set, which it clears again after the store has
//       read the buffer.
// Bound: HBM write bandwidth (302 500*m + 2 420 000 bytes per image at 550^2).
#include <stdlib.h>

#include "isb_common.cuh"

namespace isb {

constexpr int kCompThreads = 512;
constexpr int kCompWarps = kCompThreads / 32;
constexpr int kUnitPx = 4096;         // pixels per work unit (128 segments of 32)
constexpr int kYTabRows = 64;         // rows of a unit with tabulated bilinear taps (more: computed directly)
constexpr int kSmapThreads = 256;
constexpr int kSmapCtasPerSm = 4;     // persistent CTAs of the sigmoid-map kernel
constexpr int kSecondCounter = 16;    // work counter of a second compose launch (dense + packed), ints after the first

struct CompGeom {
  int units_per_img;
  int unit_px;         // pixels per work unit (<= kUnitPx; smaller when the call has few pixels in total, so every SM gets work)
  int seg_px;          // pixels per warp segment (32; 16 / 8 / 4 when m is so large that 32-pixel segment images would leave
                       // room for only a couple of warps per SM -- BASELINE configs[4], m = 1024)
  float scale_y, scale_x;
  int list_cap;        // detections per image upper bound
  int buf_bytes;       // bytes of one warp's segment buffer (multiple of 16), two per warp
  int64_t s_cap;       // floats of sigmoid-map scratch per image
};

// per-column bilinear table entry: source index pair and weight (image independent)
struct ColTap { int32_t idx; float l1; };  // idx = ix0 | (ix1 - ix0) << 16

// ---- K3a -----------------------------------------------------------------------------------
// The sigmoid-map elements of one image are one flat range [0, total_b) laid out detection after detection
// (DetRec.soff and the per-image totals come from the NMS kernel / the layout kernel).  The batch is cut into
// tickets of kSmapChunk elements, dealt round-robin to the CTAs, so the work is balanced whatever the box
// sizes (one image with large boxes does not leave the other SMs idle).  A ticket is mapped to
// (image, chunk) through the per-image ticket prefix every CTA builds in shared memory; the image's
// records sit in shared memory too, and the owning detection of an element is found by binary search.
// Dependent global round trips per CTA: totals -> records -> prototype vectors.
constexpr int kSmapChunk = 512;                            // elements per ticket
constexpr int kSmapMaxBatch = 1024;

__global__ void __launch_bounds__(kSmapThreads)
sigmoid_map_kernel(MaskParams p, const DetRec* __restrict__ rec, const int32_t* __restrict__ totals,
                   float* __restrict__ smap, int64_t s_cap, int* __restrict__ compose_counter) {
  // programmatic dependent launch: the compose kernel that follows may be scheduled as this kernel's CTAs drain and run its
  // prologue (shared-memory clears, column tap table); it waits (griddepcontrol.wait) before it touches anything written here
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");  // (launched as fast_nms_kernel's programmatic dependent in the fused path)
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // work counters of the compose kernel(s) that follow
    compose_counter[0] = 0;
    compose_counter[kSecondCounter] = 0;
  }
  if (p.n_peers > 0 && blockIdx.x == gridDim.x - 1) {  // multi-GPU: every image's count into every rank's ring (see MaskParams)
    for (int b = threadIdx.x; b < p.batch; b += kSmapThreads) {
      const unsigned long long w = ((unsigned long long)p.peer_epoch << 32) | (unsigned)p.counts[b];
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r < p.n_peers)
          asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.peer_ring[r] + p.peer_index0 + b), "l"(w) : "memory");
    }
  }
  extern __shared__ __align__(16) unsigned char s_dyn[];  // DetRec s_rec[list_cap]
  __shared__ int s_cstart[kSmapMaxBatch + 1];             // first ticket of every image
  __shared__ int s_scan[32];
  DetRec* s_rec = reinterpret_cast<DetRec*>(s_dyn);
  const int tid = threadIdx.x;
  int carry = 0;
  for (int base = 0; base < p.batch; base += kSmapThreads) {
    const int b = base + tid;
    int chunks = 0;
    if (b < p.batch) {
      const int m = p.counts ? p.counts[b] : p.m_fixed;
      if (m > 0) chunks = (totals[b] + kSmapChunk - 1) / kSmapChunk;
    }
    int tot;
    const int excl = cta_exclusive_scan(chunks, s_scan, &tot);
    if (b < p.batch) s_cstart[b] = carry + excl;
    carry += tot;
  }
  if (tid == 0) s_cstart[p.batch] = carry;
  __syncthreads();
  const int ntickets = carry;
  int cached = -1, m = 0, total = 0;
  const int per_cta = (ntickets + gridDim.x - 1) / gridDim.x;  // contiguous tickets: a CTA stays within one or two images
  const int ticket_end = min(ntickets, ((int)blockIdx.x + 1) * per_cta);
  for (int ticket = blockIdx.x * per_cta; ticket < ticket_end; ++ticket) {
    int lo = 0, hi = p.batch - 1;  // image: last b with s_cstart[b] <= ticket
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_cstart[mid] <= ticket) lo = mid; else hi = mid - 1;
    }
    const int b = lo;
    if (b != cached) {  // CTA-uniform
      __syncthreads();  // previous image's records consumed
      m = p.counts ? p.counts[b] : p.m_fixed;
      total = totals[b];
      const DetRec* rimg = rec + (size_t)b * p.det_stride;
      for (int j = tid; j < m; j += kSmapThreads) s_rec[j] = rimg[j];
      cached = b;
      __syncthreads();
    }
    const int t0 = (ticket - s_cstart[b]) * kSmapChunk;
    const float* proto_img = p.proto + (size_t)b * p.proto_h * p.proto_w * ISB_NUM_COEF;
    const float* coef_img = p.coef + (size_t)b * p.det_stride * ISB_NUM_COEF;
    float* S = smap + (size_t)b * s_cap;
    // a group of 8 lanes takes 8 consecutive elements per pass: each lane resolves (detection, pixel) of its own
    // element, the loads and partial sums are shared across the group (smap_dot8)
    const int q = tid & 7;
    int l = 0;
    DetRec r = s_rec[0];
#pragma unroll
    for (int i = 0; i < kSmapChunk / kSmapThreads; ++i) {
      const int t = t0 + i * kSmapThreads + tid;
      const bool on = t < total;  // lanes past the end recompute the last element: the shuffles need every lane
      const int tt = on ? t : total - 1;
      if (tt < r.soff || tt >= r.soff + (int)r.fw * (int)r.fh) {
        int h2 = m - 1;  // last j with soff <= t (empty footprints share an offset: take the last)
        l = 0;
        while (l < h2) {
          const int mid = (l + h2 + 1) >> 1;
          if (s_rec[mid].soff <= tt) l = mid; else h2 = mid - 1;
        }
        r = s_rec[l];
      }
      const int f = tt - r.soff, fw = r.fw;
      // f / fw without the integer division (~20 instructions): float estimate + one correction step, exact for any f
      int fy = __float2int_rz(((float)f + 0.5f) * __frcp_rn((float)fw));
      int fx = f - fy * fw;
      if (fx < 0) { --fy; fx += fw; } else if (fx >= fw) { ++fy; fx -= fw; }
      const float v = sigmoidf_ref(smap_dot8(reinterpret_cast<const float4*>(proto_img), reinterpret_cast<const float4*>(coef_img),
                                             (r.pya + fy) * p.proto_w + (r.pa + fx), l, q));
      if (on) S[t] = v;
    }
  }
}

__global__ void reset_counters_kernel(int* __restrict__ compose_counter) {
  if (threadIdx.x == 0) {
    compose_counter[0] = 0;
    compose_counter[kSecondCounter] = 0;
  }
}

// ---- K3b -----------------------------------------------------------------------------------
// zero buf[a, b) with one warp: 16-byte interior, byte edges (never touches a byte outside [a, b))
__device__ __forceinline__ void warp_clear(uint8_t* buf, int a, int b, int lane) {
  const int a16 = (a + 15) & ~15, b16 = b & ~15;
  if (a16 >= b16) {
    for (int i = a + lane; i < b; i += 32) buf[i] = 0;
    return;
  }
  if (a + lane < a16) buf[a + lane] = 0;
  if (b16 + lane < b) buf[b16 + lane] = 0;
  for (int v = a16 + lane * 16; v < b16; v += 32 * 16) *reinterpret_cast<uint4*>(buf + v) = make_uint4(0, 0, 0, 0);
}

// kMode = kCompMap is the class-map-only variant (yolact.py:162-170 needs nothing else on the host): same walk,
// no segment images, no bulk stores, one byte per pixel out.
// kMode = kCompPacked writes the SAME bool [H,W,m] stream one BIT per element instead of one byte (bit i of the image's
// packed stream = byte i of the dense stream, little-endian within 32-bit words) plus the dense int64 arg-max: the
// transport format of isb_yolact_detect_host, which expands it into the caller's host buffer on host threads -- an
// eighth of the bytes cross PCIe.  A 32-pixel segment is exactly m 32-bit words (32 * m bits, word-aligned because
// segments start at multiples of 32 pixels): lanes OR their bits into the warp's m-word buffer, which then leaves as
// plain coalesced stores.
enum { kCompMap = 0, kCompDense = 1, kCompPacked = 2 };

// kBig: the one-CTA-per-SM configuration of very large detection counts (BASELINE configs[4]: m = 1024, 8-pixel segments,
// ~10 warps per SM).  With so few warps the kernel is bound by its serial memory round trips, so this instantiation may
// use 128 registers and (a) loads the records of up to four list-building rounds before the first barrier instead of one
// round trip + three barriers per round, (b) keeps the sigmoid-map taps of TWO covering detections in flight per lane.
template <int kMode, bool kBig>
__global__ void __launch_bounds__(kCompThreads, kBig ? 1 : 2)
mask_compose_kernel(MaskParams p, CompGeom g, const DetRec* __restrict__ rec, const float* __restrict__ smap,
                    int* __restrict__ work_counter) {
  constexpr bool kMasks = kMode == kCompDense;
  constexpr bool kPacked = kMode == kCompPacked;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ int s_unit;
  __shared__ int s_wcount[kCompWarps];
  __shared__ int s_nlist;
  __shared__ ColTap s_ytab[kYTabRows];

  const int H = p.out_h, W = p.out_w;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nwarps = blockDim.x >> 5;
  // layout: [nwarps][2] segment buffers | column taps | listed detection records
  uint8_t* wbuf = smem_raw + (size_t)warp * 2 * g.buf_bytes;
  const size_t off_tab = (size_t)nwarps * 2 * g.buf_bytes;
  ColTap* s_xtab = reinterpret_cast<ColTap*>(smem_raw + off_tab);
  DetRec* s_det = reinterpret_cast<DetRec*>(smem_raw + off_tab + (((size_t)W * sizeof(ColTap) + 31) & ~(size_t)31));

  const int total_units = p.batch * g.units_per_img;
  const int HW = H * W;

  // ---- once per CTA: clean segment buffers, per-column bilinear taps -----------------------------
  if (kMasks)  // (the packed variant clears its m words per segment)
    for (int i = tid; i < nwarps * 2 * g.buf_bytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem_raw)[i] = make_uint4(0, 0, 0, 0);
  for (int x = tid; x < W; x += blockDim.x) {
    int ix0, ix1;
    float l0, l1;
    bilinear_src(x, g.scale_x, p.proto_w, ix0, ix1, l0, l1);
    ColTap t;
    t.idx = ix0 | ((ix1 - ix0) << 16);
    t.l1 = l1;
    s_xtab[x] = t;
  }
  int dlo0 = 0, dhi0 = 0, dlo1 = 0, dhi1 = 0;  // this warp's dirty byte range in each of its buffers
  unsigned seg_no = 0;
  // everything above is independent of the kernels before this one; counts, records, sigmoid maps and the work counter are not
  asm volatile("griddepcontrol.wait;" ::: "memory");

  int next_unit = 0;  // kBig: the ticket of the following unit is fetched while this one is processed (units are short there)
  if (kBig && tid == 0) next_unit = atomicAdd(work_counter, 1);
  while (true) {
    __syncthreads();  // everyone is done with the previous unit's list (and s_unit)
    if (tid == 0) s_unit = kBig ? next_unit : atomicAdd(work_counter, 1);
    __syncthreads();
    const int unit = s_unit;
    if (unit >= total_units) break;
    if (kBig && tid == 0) next_unit = atomicAdd(work_counter, 1);
    const int b = unit / g.units_per_img;
    const int u = unit - b * g.units_per_img;
    const int m = p.counts ? p.counts[b] : p.m_fixed;
    if (m <= 0) {  // nothing to write for this image (the reference returns None); its class map is background
      if (p.class_map) {
        const int px0 = u * g.unit_px, px1 = min(px0 + g.unit_px, HW);
        for (int q = px0 + tid; q < px1; q += blockDim.x) p.class_map[(size_t)b * HW + q] = 0;
      }
      continue;
    }
    const int px0 = u * g.unit_px;
    const int px1 = min(px0 + g.unit_px, HW);
    const int segpx = g.seg_px;
    const int ya = px0 / W, yb = (px1 - 1) / W;  // rows the unit touches
    uint8_t* out_img = kMasks ? p.masks + (size_t)b * p.mask_slot_bytes : nullptr;
    uint32_t* pk_img = kPacked ? reinterpret_cast<uint32_t*>(p.packed + (size_t)b * p.packed_slot_bytes) : nullptr;
    int64_t* arg_img = (kMasks || kPacked) ? p.masks_arg + (size_t)b * HW : nullptr;
    uint8_t* cmap_img = p.class_map ? p.class_map + (size_t)b * HW : nullptr;
    const int64_t* cls_img = p.class_ids + (size_t)b * p.class_stride;
    const float* S_img = smap + (size_t)b * g.s_cap;

    // ---- per-row bilinear taps of the unit's rows, detections that meet those rows (ascending index) ----
    const bool ytab_ok = yb - ya < kYTabRows;
    if (ytab_ok && tid <= yb - ya) {
      int iy0, iy1;
      float l0, l1;
      bilinear_src(ya + tid, g.scale_y, p.proto_h, iy0, iy1, l0, l1);
      ColTap t;
      t.idx = iy0 | ((iy1 - iy0) << 16);
      t.l1 = l1;
      s_ytab[tid] = t;
    }
    if (tid == 0) s_nlist = 0;
    __syncthreads();
    if constexpr (kBig) {
      // four rounds per sweep: all record loads of the sweep are in flight together, one count exchange for the four
      for (int base = 0; base < m; base += 4 * blockDim.x) {
        int4 bb[4];
        unsigned bal[4];
        bool hit[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int j = base + q * blockDim.x + tid;
          bb[q] = make_int4(0, 0, 0, 0);
          if (j < m) bb[q] = __ldg(reinterpret_cast<const int4*>(&rec[(size_t)b * p.det_stride + j]));  // xlo, xhi, ylo, yhi
        }
        int packed = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int j = base + q * blockDim.x + tid;
          hit[q] = j < m && bb[q].z <= yb && bb[q].w > ya && bb[q].x < bb[q].y;
          bal[q] = __ballot_sync(0xffffffffu, hit[q]);
          packed |= __popc(bal[q]) << (8 * q);  // <= 32 per round and warp
        }
        if (lane == 0) s_wcount[warp] = packed;
        __syncthreads();
        int before[4] = {0, 0, 0, 0}, tot[4] = {0, 0, 0, 0};
        for (int w = 0; w < nwarps; ++w) {
          const int c = s_wcount[w];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int cq = (c >> (8 * q)) & 0xff;
            if (w < warp) before[q] += cq;
            tot[q] += cq;
          }
        }
        int off = s_nlist;  // ascending detection index: round-major, then warp, then lane
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (hit[q]) {
            const int j = base + q * blockDim.x + tid;
            s_det[off + before[q] + __popc(bal[q] & ((1u << lane) - 1u))] = rec[(size_t)b * p.det_stride + j];
          }
          off += tot[q];
        }
        __syncthreads();
        if (tid == 0) s_nlist = off;
        __syncthreads();
      }
    } else {
    for (int base = 0; base < m; base += blockDim.x) {
      const int j = base + tid;
      bool hit = false;
      DetRec r;
      if (j < m) {
        r = rec[(size_t)b * p.det_stride + j];
        hit = r.ylo <= yb && r.yhi > ya && r.xlo < r.xhi;
      }
      const unsigned bal = __ballot_sync(0xffffffffu, hit);
      if (lane == 0) s_wcount[warp] = __popc(bal);
      __syncthreads();
      int off = s_nlist, tot = 0;
      for (int w = 0; w < nwarps; ++w) {
        const int c = s_wcount[w];
        if (w < warp) off += c;
        tot += c;
      }
      if (hit) s_det[off + __popc(bal & ((1u << lane) - 1u))] = r;
      __syncthreads();
      if (tid == 0) s_nlist += tot;
      __syncthreads();
    }
    }
    const int nlist = s_nlist;

    // ---- segments of 32 pixels: every warp runs its own, independently of the other warps ---------------
    int xq = (px0 - ya * W) + warp * segpx + lane, yq = ya;  // this lane's pixel in the warp's first segment
    while (xq >= W) { xq -= W; ++yq; }
    for (int q0 = px0 + warp * segpx; q0 < px1; q0 += nwarps * segpx, ++seg_no) {
      const int nt = min(segpx, px1 - q0);
      const int par = seg_no & 1;
      uint8_t* buf = wbuf + (par ? g.buf_bytes : 0);
      uint8_t* g0 = kMasks ? out_img + (int64_t)q0 * m : nullptr;
      const int shift = kMasks ? (int)(reinterpret_cast<uintptr_t>(g0) & 15) : 0;
      if (kMasks) {
        // the bulk store issued from this buffer two segments ago must have finished reading it
        if (lane == 0) bulk_wait_read1();
        __syncwarp();
        const int dlo = par ? dlo1 : dlo0, dhi = par ? dhi1 : dhi0;
        if (dhi > dlo) {
          warp_clear(buf, dlo, dhi, lane);
          __syncwarp();
        }
      }
      uint32_t* wbits = reinterpret_cast<uint32_t*>(wbuf);  // packed mode: the warp's m words
      if (kPacked) {
        __syncwarp();  // every lane has stored the previous segment's words
        for (int w = lane; w < m; w += 32) wbits[w] = 0u;
        __syncwarp();
      }
      const bool active = lane < nt;
      const int y = yq, x = xq;
      const int wy0 = __shfl_sync(0xffffffffu, y, 0), wy1 = __shfl_sync(0xffffffffu, y, nt - 1);
      int wxa = 0, wxb = W;
      if (wy0 == wy1) { wxa = __shfl_sync(0xffffffffu, x, 0); wxb = wxa + nt; }
      int iy0, iy1;
      float ly0, ly1;
      if (ytab_ok) {
        const ColTap ty = s_ytab[min(y, yb) - ya];
        iy0 = ty.idx & 0xffff; iy1 = iy0 + (ty.idx >> 16);
        ly1 = ty.l1; ly0 = 1.0f - ly1;
      } else {
        bilinear_src(min(y, H - 1), g.scale_y, p.proto_h, iy0, iy1, ly0, ly1);
      }
      const ColTap tb = s_xtab[x];
      const int ix0 = tb.idx & 0xffff, ix1 = ix0 + (tb.idx >> 16);
      const float lx1 = tb.l1, lx0 = 1.0f - lx1;
      uint8_t* mypix = kMasks ? buf + shift + lane * m : nullptr;
      float best = 0.0f;
      int arg = 0;
      bool anyset = false;
      for (int base = 0; base < nlist; base += 32) {
        bool hit = false;
        if (base + lane < nlist) {
          const int4 bb = *reinterpret_cast<const int4*>(&s_det[base + lane]);  // xlo, xhi, ylo, yhi
          hit = bb.z <= wy1 && bb.w > wy0 && bb.x < wxb && bb.y > wxa;
        }
        unsigned bal = __ballot_sync(0xffffffffu, hit);
        if constexpr (kBig) {
          // two listed detections per step: eight sigmoid-map taps in flight per lane instead of four
          auto taps = [&](const DetRec& r, float& p00, float& p01, float& p10, float& p11) {
            const float* S = S_img + r.soff;
            const int fw = r.fw;
            const int r0 = (iy0 - r.pya) * fw - r.pa, r1 = (iy1 - r.pya) * fw - r.pa;
            p00 = __ldg(S + r0 + ix0); p01 = __ldg(S + r0 + ix1);
            p10 = __ldg(S + r1 + ix0); p11 = __ldg(S + r1 + ix1);
          };
          auto take = [&](int jdet, float p00, float p01, float p10, float p11) {
            const float top = lx0 * p00 + lx1 * p01;
            const float bot = lx0 * p10 + lx1 * p11;
            const float v = ly0 * top + ly1 * bot;
            if (kMasks && v > 0.5f) { mypix[jdet] = 1; anyset = true; }
            if (v > best) { best = v; arg = jdet; }
          };
          while (bal) {
            const int k = __ffs(bal) - 1;
            bal &= bal - 1;
            const bool two = bal != 0u;  // warp-uniform
            const int k2 = two ? __ffs(bal) - 1 : k;
            if (two) bal &= bal - 1;
            const DetRec r = s_det[base + k], r2 = s_det[base + k2];
            const bool in0 = active && y >= r.ylo && y < r.yhi && x >= r.xlo && x < r.xhi;
            const bool in1 = two && active && y >= r2.ylo && y < r2.yhi && x >= r2.xlo && x < r2.xhi;
            float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f, b00 = 0.f, b01 = 0.f, b10 = 0.f, b11 = 0.f;
            if (in0) taps(r, a00, a01, a10, a11);
            if (in1) taps(r2, b00, b01, b10, b11);
            if (in0) take(r.j, a00, a01, a10, a11);
            if (in1) take(r2.j, b00, b01, b10, b11);  // ascending index: first max wins
          }
        } else {
        while (bal) {
          const int k = __ffs(bal) - 1;
          bal &= bal - 1;
          const DetRec r = s_det[base + k];
          if (active && y >= r.ylo && y < r.yhi && x >= r.xlo && x < r.xhi) {
            const float* S = S_img + r.soff;
            const int fw = r.fw;
            const int r0 = (iy0 - r.pya) * fw - r.pa, r1 = (iy1 - r.pya) * fw - r.pa;
            const float top = lx0 * __ldg(S + r0 + ix0) + lx1 * __ldg(S + r0 + ix1);
            const float bot = lx0 * __ldg(S + r1 + ix0) + lx1 * __ldg(S + r1 + ix1);
            const float v = ly0 * top + ly1 * bot;
            if (kMasks && v > 0.5f) { mypix[r.j] = 1; anyset = true; }
            if (kPacked && v > 0.5f) {
              const int bit = lane * m + r.j;
              atomicOr(&wbits[bit >> 5], 1u << (bit & 31));
            }
            if (v > best) { best = v; arg = r.j; }
          }
        }
        }
      }
      if (cmap_img && active) cmap_img[q0 + lane] = best > 0.5f ? (uint8_t)(__ldg(cls_img + arg) + 1) : (uint8_t)0;
      if (kMasks) {
        if (active) arg_img[q0 + lane] = (int64_t)arg;
        {  // remember what this warp dirtied
          const unsigned sb = __ballot_sync(0xffffffffu, anyset);
          int dlo = 0, dhi = 0;
          if (sb) {
            dlo = shift + (__ffs(sb) - 1) * m;
            dhi = shift + (32 - __clz(sb)) * m;
          }
          if (par) { dlo1 = dlo; dhi1 = dhi; } else { dlo0 = dlo; dhi0 = dhi; }
        }
        fence_proxy_async();  // this lane's shared-memory writes -> visible to the bulk store
        __syncwarp();
        // ---- the segment leaves as one bulk store (16-byte aligned interior) + < 16-byte edges ---------
        const int nbytes = nt * m;
        const int head = min((16 - shift) & 15, nbytes);
        const int body = (nbytes - head) & ~15;
        const int tail = nbytes - head - body;
        if (lane == 0) {
          if (body > 0) bulk_s2g(g0 + head, buf + shift + head, (uint32_t)body);
          bulk_commit();  // one (possibly empty) group per segment keeps the wait_group arithmetic exact
        }
        if (lane < head) g0[lane] = buf[shift + lane];
        if (lane >= 16 && lane - 16 < tail) g0[head + body + (lane - 16)] = buf[shift + head + body + (lane - 16)];
      }
      if (kPacked) {
        if (active && p.skip_dense) arg_img[q0 + lane] = (int64_t)arg;  // (a dense pass before this one wrote it already)
        __syncwarp();  // all bits of the segment are in shared memory
        const int nwords = (nt * m + 31) >> 5;  // nt < 32 only in the image's last segment: the tail word is zero-padded
        uint32_t* gw = pk_img + (size_t)(q0 >> 5) * m;
        for (int w = lane; w < nwords; w += 32) gw[w] = wbits[w];
      }
      // this lane's pixel in the warp's next segment
      xq += nwarps * segpx;
      while (xq >= W) { xq -= W; ++yq; }
    }
  }
  if (kMasks && lane == 0) bulk_wait_read0();  // shared memory must outlive every bulk store that reads it
}

// ---- host side ----------------------------------------------------------------------------
static constexpr int kCompSmemTwo = 111 * 1024;   // two CTAs per SM
static constexpr int kCompSmemOne = 224 * 1024;   // one CTA per SM (very large detection counts)

int configure_masks(isb_handle* h) {
  ISB_CUDA(h, cudaFuncSetAttribute(mask_compose_kernel<kCompDense, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCompSmemOne));
  ISB_CUDA(h, cudaFuncSetAttribute(mask_compose_kernel<kCompDense, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCompSmemOne));
  ISB_CUDA(h, cudaFuncSetAttribute(mask_compose_kernel<kCompMap, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCompSmemOne));
  ISB_CUDA(h, cudaFuncSetAttribute(mask_compose_kernel<kCompPacked, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kCompSmemOne));
  return ISB_OK;
}

static int ensure_smap(isb_handle* h, size_t floats) {
  if (floats <= h->smap_floats) return ISB_OK;
  ISB_CUDA(h, cudaDeviceSynchronize());
  cudaFree(h->smap);
  h->smap = nullptr;
  h->smap_floats = 0;
  ISB_CUDA(h, cudaMalloc((void**)&h->smap, floats * sizeof(float)));
  h->smap_floats = floats;
  return ISB_OK;
}

int ensure_mask_scratch(isb_handle* h, int proto_h, int proto_w, int m_cap, int64_t images, int64_t* s_cap) {
  // worst case per image: every box covers the whole prototype map.  Sized by the CALL (its detection bound and image
  // count), not by the handle's limits: 7.6 MB per image at m = 100, 138 x 138.
  *s_cap = (int64_t)m_cap * proto_h * proto_w;
  return ensure_smap(h, (size_t)images * (size_t)*s_cap);
}

bool masks_tc_smap_forced() {
  static const bool v = getenv("ISB_TC_SMAP") != nullptr;
  return v;
}

int launch_masks(isb_handle* h, const MaskParams& p, cudaStream_t s, int which) {
  if (p.proto_w >= 32768 || p.proto_h >= 32768) ISB_FAIL(h, ISB_ERR_LIMIT, "mask assembly supports prototype sides < 32768");
  if ((int64_t)p.out_h * p.out_w >= (1ll << 31) / 8) ISB_FAIL(h, ISB_ERR_LIMIT, "mask assembly: output image too large");
  if (p.det_stride > h->lim.max_det) ISB_FAIL(h, ISB_ERR_LIMIT, "mask assembly: %d detections > max_det %d", p.det_stride, h->lim.max_det);
  CompGeom g;
  const int64_t hw = (int64_t)p.out_h * p.out_w;
  // work units: kUnitPx pixels, halved while the call has fewer than ~4 units per resident CTA (one image with very many
  // detections -- configs[4] -- would otherwise keep half the SMs idle: 74 units for 148 CTAs)
  g.unit_px = kUnitPx;
  while (g.unit_px > 512 && (int64_t)p.batch * ((hw + g.unit_px - 1) / g.unit_px) < 8ll * h->num_sms) g.unit_px >>= 1;
  g.units_per_img = (int)((hw + g.unit_px - 1) / g.unit_px);
  g.seg_px = 32;
  g.scale_y = (float)p.proto_h / (float)p.out_h;
  g.scale_x = (float)p.proto_w / (float)p.out_w;
  g.list_cap = p.m_fixed >= 0 ? (p.m_fixed > 0 ? p.m_fixed : 1) : (p.m_cap > 0 ? p.m_cap : p.det_stride);
  const size_t fixed = (((size_t)p.out_w * sizeof(ColTap) + 31) & ~(size_t)31) + (size_t)g.list_cap * sizeof(DetRec);
  if (p.class_map && !p.class_ids) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask assembly: class_map needs class_ids");
  if (p.map_only && !p.class_map) ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask assembly: map_only without class_map");
  if (p.packed && (p.packed_slot_bytes < 4 * ((hw * g.list_cap + 31) / 32) || (p.packed_slot_bytes & 15)))
    ISB_FAIL(h, ISB_ERR_BAD_ARG, "mask assembly: packed_slot_bytes must be >= 4*ceil(H*W*max_det/32) and a multiple of 16");
  const int64_t ws_images = h->ws_images > h->ws_b0 + p.batch ? h->ws_images : h->ws_b0 + p.batch;
  int rc = ensure_mask_scratch(h, p.proto_h, p.proto_w, g.list_cap, ws_images, &g.s_cap);
  if (rc != ISB_OK) return rc;
  const int64_t total_units = (int64_t)p.batch * g.units_per_img;
  if (total_units == 0) return ISB_OK;
  DetRec* rec = h->det_rec + (size_t)h->ws_b0 * h->lim.max_det;
  float* smap = h->smap + (size_t)h->ws_b0 * g.s_cap;
  if (p.batch > kSmapMaxBatch) ISB_FAIL(h, ISB_ERR_LIMIT, "mask assembly supports at most %d images per call", kSmapMaxBatch);
  int* counters = h->work_counter + h->ws_lane;
  static const bool tc_smap = getenv("ISB_TC_SMAP") != nullptr;  // tensor-core product instead of K3a (measurement switch)
  if (!(which & 1)) {
    // compose only: the sigmoid maps were launched separately
  } else if (tc_smap && g.list_cap <= kTcBatchMaxDet) {
    reset_counters_kernel<<<1, 32, 0, s>>>(counters);
    ISB_LAUNCH_CHECK(h, "reset_counters_kernel");
    TcBatchArgs a;
    memset(&a, 0, sizeof(a));
    a.proto = p.proto; a.coef = p.coef; a.coef_rows = p.det_stride; a.counts = p.counts; a.m_fixed = p.m_fixed;
    a.batch = p.batch; a.npix = p.proto_h * p.proto_w; a.proto_w = p.proto_w; a.mode = 2;
    a.rec = rec; a.rec_stride = p.det_stride; a.smap = smap; a.s_cap = g.s_cap;
    if ((rc = launch_mask_logits_batch(h, a, s)) != ISB_OK) return rc;
  } else {
    cudaLaunchConfig_t lc;
    memset(&lc, 0, sizeof(lc));
    lc.gridDim = dim3((unsigned)(h->num_sms * kSmapCtasPerSm));
    lc.blockDim = dim3(kSmapThreads);
    lc.dynamicSmemBytes = (size_t)g.list_cap * sizeof(DetRec);
    lc.stream = s;
    cudaLaunchAttribute attr;
    memset(&attr, 0, sizeof(attr));
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    static const bool no_pdl = getenv("ISB_NO_PDL") != nullptr;
    lc.attrs = &attr;
    lc.numAttrs = (h->smap_follows_nms && !no_pdl) ? 1 : 0;
    const cudaError_t le = cudaLaunchKernelEx(&lc, sigmoid_map_kernel, p, (const DetRec*)rec, (const int32_t*)(h->smap_total + h->ws_b0),
                                              smap, g.s_cap, counters);
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "sigmoid_map_kernel launch: %s", cudaGetErrorString(le));
    ISB_LAUNCH_CHECK(h, "sigmoid_map_kernel");
  }
  if (h->prof_mid) ISB_CUDA(h, cudaEventRecord(h->prof_mid, s));
  if (!(which & 2)) return ISB_OK;
  // one compose launch per output format: class map only | dense bytes | packed bits (dense + packed when both are asked for)
  const int modes[2] = {p.map_only ? kCompMap : ((p.packed && p.skip_dense) ? kCompPacked : kCompDense),
                        (!p.map_only && p.packed && !p.skip_dense) ? kCompPacked : -1};
  for (int pass = 0; pass < 2; ++pass) {
    const int mode = modes[pass];
    if (mode < 0) break;
    // per-warp staging: dense = two segment images of 32 pixels x m bytes (+ alignment shift); packed = m words (split
    // over the two "buffers"); the class-map-only variant has none
    size_t buf = 0;
    size_t budget = kCompSmemTwo;
    int nwarps = kCompWarps;
    g.seg_px = 32;
    if (mode == kCompPacked) buf = ((size_t)2 * g.list_cap + 15) & ~(size_t)15;
    if (mode == kCompDense) {
      // as many warps as fit: two CTAs per SM if that gives >= 8 warps, else one; if even that leaves fewer than 8
      // warps, shorter segments (the segment image is seg_px x m bytes)
      for (int seg = 32; seg >= 4; seg >>= 1) {
        g.seg_px = seg;
        buf = ((size_t)seg * g.list_cap + 16 + 15) & ~(size_t)15;
        budget = kCompSmemTwo;
        nwarps = fixed < budget ? (int)((budget - fixed) / (2 * buf)) : 0;
        if (nwarps >= 8) break;
        budget = kCompSmemOne;
        nwarps = fixed < budget ? (int)((budget - fixed) / (2 * buf)) : 0;
        if (nwarps >= 8) break;
      }
    } else if (mode == kCompPacked) {
      nwarps = fixed < budget ? (int)((budget - fixed) / (2 * buf)) : 0;
      if (nwarps < 8) {
        budget = kCompSmemOne;
        nwarps = fixed < budget ? (int)((budget - fixed) / (2 * buf)) : 0;
      }
    } else if (fixed > budget) {
      budget = kCompSmemOne;
    }
    if (nwarps > kCompWarps) nwarps = kCompWarps;
    if (nwarps < 1 || fixed > budget)
      ISB_FAIL(h, ISB_ERR_LIMIT, "mask compose: %d detections x width %d exceed shared memory", g.list_cap, p.out_w);
    g.buf_bytes = (int)buf;
    const size_t smem = fixed + (size_t)nwarps * 2 * buf;
    const int per_sm = budget == (size_t)kCompSmemTwo ? 2 : 1;
    int64_t grid = (int64_t)h->num_sms * per_sm;
    if (grid > total_units) grid = total_units;
    int* ctr = counters + (pass ? kSecondCounter : 0);
    cudaLaunchConfig_t lc;
    memset(&lc, 0, sizeof(lc));
    lc.gridDim = dim3((unsigned)grid);
    lc.blockDim = dim3((unsigned)(nwarps * 32));
    lc.dynamicSmemBytes = smem;
    lc.stream = s;
    cudaLaunchAttribute attr;
    memset(&attr, 0, sizeof(attr));
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    static const bool no_pdl = getenv("ISB_NO_PDL") != nullptr;
    lc.attrs = &attr;
    // only right behind the sigmoid-map kernel of this call (the fused path says so; isb_mask_assemble launches both here)
    const bool behind_smap = h->compose_follows_smap || ((which & 3) == 3 && !h->prof_mid && !(tc_smap && g.list_cap <= kTcBatchMaxDet));
    lc.numAttrs = (pass == 0 && behind_smap && !no_pdl) ? 1 : 0;
    cudaError_t le;
    if (mode == kCompMap)
      le = cudaLaunchKernelEx(&lc, mask_compose_kernel<kCompMap, false>, p, g, (const DetRec*)rec, (const float*)smap, ctr);
    else if (mode == kCompDense && per_sm == 1)  // very large detection counts: the latency-oriented instantiation
      le = cudaLaunchKernelEx(&lc, mask_compose_kernel<kCompDense, true>, p, g, (const DetRec*)rec, (const float*)smap, ctr);
    else if (mode == kCompDense)
      le = cudaLaunchKernelEx(&lc, mask_compose_kernel<kCompDense, false>, p, g, (const DetRec*)rec, (const float*)smap, ctr);
    else
      le = cudaLaunchKernelEx(&lc, mask_compose_kernel<kCompPacked, false>, p, g, (const DetRec*)rec, (const float*)smap, ctr);
    if (le != cudaSuccess) ISB_FAIL(h, ISB_ERR_CUDA, "mask_compose_kernel launch: %s", cudaGetErrorString(le));
    ISB_LAUNCH_CHECK(h, "mask_compose_kernel");
  }
  return ISB_OK;
}

}  // namespace isb
