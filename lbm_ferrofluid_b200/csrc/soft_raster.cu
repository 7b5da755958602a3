// Soft rasteriser, forward pass (second half): the CUDA kernel behind the frames the 3-D demos
// write (src/LBM/utils/io.py:50-84 -> renderutils/softrenderer.py:266-467 -> renderutils/cuda/
// soft_rasterize_cuda_kernel.cu:236-475, "Soft Rasterizer", Liu et al. 2019).
//
// The reference launches one thread per pixel and every thread walks ALL faces: 512 x 512 pixels x ~10^5..10^6
// marching-cubes triangles per frame, of which a pixel is touched by a handful (a face is skipped unless the pixel
// lies inside its bounding box grown by sqrt(threshold), soft_rasterize_cuda_kernel.cu:372).  Here:
//   1. k_face_setup     per face: barycentric matrix, Gram matrix, obtuse-corner flag (:236-302) + the grown bbox;
//   2. k_group_box/k_bin per 8 x 8 pixel tile, warps scan the face boxes IN FACE ORDER (first the union boxes of
//                       groups of 32 consecutive faces, then the faces of the groups that touch the tile) and compacts
//                       the hits (ballot + popc): ordered per-tile lists, no atomics, deterministic;
//   3. k_raster_tile    one block (two warps) per tile, one thread per pixel: the tile's faces are staged through shared memory
//                       in chunks and every pixel folds them in ascending face order -- the same aggregation order
//                       as the reference's loop over all faces, so the running softmax rescaling sees the same
//                       sequence of faces (those the reference skips by its bbox test contribute nothing there either).
// Every option of the reference forward is kept: dist hard / barycentric / euclidean, rgb hard / softmax, alpha
// hard / sum / prod, surface / vertex textures, near / far, double_side.  fp32 like the reference, and like the
// reference (--use_fast_math) the per-pixel code uses approximate division / exp (the comparison tolerance against
// the IEEE numpy oracle and the reference kernel is in the tests).
#include "capi.cuh"

namespace lbmff {

constexpr int kTile = 8;         // pixels per tile edge: small tiles = short lists and thousands of blocks to balance
constexpr int kFW = 8, kFH = 4;  // a warp's pixel footprint inside the tile
constexpr int kSeg = 4;          // binning warps per tile: each scans a quarter of the faces
constexpr int kChunk = 64;       // faces staged per shared-memory chunk
constexpr int kInfo = 24;        // per-face record: 9 inverse, 9 Gram, 3 obtuse flags, 3 padding

struct RasterParams {
  int B, F, is, T, R;            // batch, faces, image size, texels per face, texture resolution
  float near_, far_, eps, sigma, dist_eps, gamma;
  int dist, rgb, alpha, sample, double_side;
  int tiles;                     // tiles per image edge
};

__device__ __forceinline__ float pix_coord(int i, int is) { return (2.0f * i + 1.0f - is) / is; }

// ---- 1. per-face setup ---------------------------------------------------------------------------------------
// info: [0..8] rows of the inverse of [[x0 x1 x2][y0 y1 y2][1 1 1]] (barycentric weights = M (x, y, 1)),
//       [9..17] Gram matrix (x_j x_k + y_j y_k + 1), [18..20] 1 at the first obtuse corner.  box: xmin xmax ymin ymax
//       of the vertices, grown by sqrt(threshold) as the pixel test does (kernel.cu:48-53,372).
__global__ void __launch_bounds__(256) k_face_setup(const float* __restrict__ faces, float* __restrict__ info,
                                                    float4* __restrict__ box, int n, float grow) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* f = faces + (long long)i * 9;
  const float x0 = f[0], y0 = f[1], x1 = f[3], y1 = f[4], x2 = f[6], y2 = f[7];
  float det = x2 * (y0 - y1) + x0 * (y1 - y2) + x1 * (y2 - y0);
  det = det > 0 ? fmaxf(det, 1e-10f) : fminf(det, -1e-10f);
  const float adj[9] = {y1 - y2, x2 - x1, x1 * y2 - x2 * y1,
                        y2 - y0, x0 - x2, x2 * y0 - x0 * y2,
                        y0 - y1, x1 - x0, x0 * y1 - x1 * y0};
  float* o = info + (long long)i * kInfo;
#pragma unroll
  for (int k = 0; k < 9; ++k) o[k] = adj[k] / det;
  const float px[3] = {x0, x1, x2}, py[3] = {y0, y1, y2};
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int k = 0; k < 3; ++k) o[9 + 3 * j + k] = px[j] * px[k] + py[j] * py[k] + 1.0f;
  float obt[3] = {0.f, 0.f, 0.f};
  for (int k = 0; k < 3; ++k) {
    const int k1 = (k + 1) % 3, k2 = (k + 2) % 3;
    if ((px[k1] - px[k]) * (px[k2] - px[k]) + (py[k1] - py[k]) * (py[k2] - py[k]) < 0) {
      obt[k] = 1.f;
      break;
    }
  }
  o[18] = obt[0]; o[19] = obt[1]; o[20] = obt[2]; o[21] = o[22] = o[23] = 0.f;
  box[i] = make_float4(fminf(fminf(x0, x1), x2) - grow, fmaxf(fmaxf(x0, x1), x2) + grow,
                       fminf(fminf(y0, y1), y2) - grow, fmaxf(fmaxf(y0, y1), y2) + grow);
}

// ---- 2. binning ------------------------------------------------------------------------------------------------
// A pixel (xp, yp) keeps a face iff xmin - g <= xp <= xmax + g and ymin - g <= yp <= ymax + g; xp, yp are monotone
// in the pixel index, so testing the tile's extreme pixel centres is a superset of its pixels' tests.
struct TileSpan {
  float xlo, xhi, ylo, yhi;
};
__device__ __forceinline__ TileSpan tile_span(int tx, int ty, int is) {
  TileSpan s;
  const int x0 = tx * kTile, x1 = min(x0 + kTile, is) - 1;
  const int r0 = ty * kTile, r1 = min(r0 + kTile, is) - 1;  // pixel rows; row r is y index is - 1 - r
  s.xlo = pix_coord(x0, is); s.xhi = pix_coord(x1, is);
  s.ylo = pix_coord(is - 1 - r1, is); s.yhi = pix_coord(is - 1 - r0, is);
  return s;
}
__device__ __forceinline__ bool touches(const float4 b, const TileSpan& s) {
  return !(s.xlo > b.y || s.xhi < b.x || s.ylo > b.w || s.yhi < b.z);
}

// union box of each group of 32 consecutive faces (one warp per group).  Marching-cubes faces arrive in cell order,
// so a group is a compact patch of the surface and its box is tight: a tile warp rejects 32 faces with one test.
__global__ void __launch_bounds__(256) k_group_box(const float4* __restrict__ box, float4* __restrict__ gbox, int F,
                                                   int groups_per_item, int ngroups_total) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= ngroups_total) return;
  const int bn = w / groups_per_item, g = w - bn * groups_per_item, f = g * 32 + lane;
  float4 b = make_float4(3.0e38f, -3.0e38f, 3.0e38f, -3.0e38f);  // empty
  if (f < F) b = __ldg(box + (long long)bn * F + f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    b.x = fminf(b.x, __shfl_xor_sync(0xffffffffu, b.x, o));
    b.y = fmaxf(b.y, __shfl_xor_sync(0xffffffffu, b.y, o));
    b.z = fminf(b.z, __shfl_xor_sync(0xffffffffu, b.z, o));
    b.w = fmaxf(b.w, __shfl_xor_sync(0xffffffffu, b.w, o));
  }
  if (lane == 0) gbox[w] = b;
}

// kSeg warps per (batch item, tile), each on its share of the faces: scans the group boxes 32 at a time, then the 32 faces of every group that touches
// the tile, all in ascending face order.  FILL = false counts, FILL = true writes the ordered face list.
template <bool FILL>
__global__ void __launch_bounds__(256) k_bin(const float4* __restrict__ box, const float4* __restrict__ gbox, int F,
                                             int groups_per_item, int is, int tiles, int ntile_total,
                                             int* __restrict__ count_or_offset, int* __restrict__ list) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= ntile_total * kSeg) return;
  const int tile = w / kSeg, seg = w - tile * kSeg;
  const int per_img = tiles * tiles, bn = tile / per_img, t = tile - bn * per_img;
  const TileSpan s = tile_span(t % tiles, t / tiles, is);
  const float4* b = box + (long long)bn * F;
  const float4* gb = gbox + (long long)bn * groups_per_item;
  // segment seg owns a contiguous range of face groups (multiples of 32 groups); list order = (tile, seg, face)
  const int per_seg = ((groups_per_item + kSeg - 1) / kSeg + 31) / 32 * 32;
  const int g_beg = seg * per_seg, g_end = min(groups_per_item, g_beg + per_seg);
  int n = 0;
  const int base = FILL ? count_or_offset[w] : 0;
  for (int g0 = g_beg; g0 < g_end; g0 += 32) {
    const int g = g0 + lane;
    unsigned gm = __ballot_sync(0xffffffffu, g < g_end && touches(__ldg(gb + g), s));
    while (gm) {
      const int bit = __ffs(gm) - 1;
      gm &= gm - 1;
      const int f = (g0 + bit) * 32 + lane;
      const bool hit = f < F && touches(__ldg(b + f), s);
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (FILL && hit) list[base + n + __popc(m & ((1u << lane) - 1u))] = f;
      n += __popc(m);
    }
  }
  if (!FILL && lane == 0) count_or_offset[w] = n;
}

// exclusive scan by one block, in place; counts[n] receives the total
__global__ void __launch_bounds__(1024) k_scan_exclusive(int* __restrict__ counts, int n) {
  __shared__ int warp_tot[32];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < n ? counts[i] : 0;
    int s = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += t;
    }
    if (lane == 31) warp_tot[warp] = s;
    __syncthreads();
    if (warp == 0) {
      int x = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += t;
      }
      warp_tot[lane] = x;
    }
    __syncthreads();
    const int before = carry + (warp ? warp_tot[warp - 1] : 0) + (s - v);
    if (i < n) counts[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[n] = carry;
}

// ---- 3. per-pixel aggregation ---------------------------------------------------------------------------------
// The reference builds its kernel with --use_fast_math (soft_rasterize.py:18-23): approximate division and exp are
// what it runs; the same intrinsics here cut the per-visit instruction count by about a third.
__device__ __forceinline__ float fdiv(float a, float b) { return __fdividef(a, b); }

struct FaceRec {              // one staged face
  float v[9];                 // x0 y0 z0 x1 y1 z1 x2 y2 z2 (projected: x, y in [-1, 1], z = camera depth)
  float info[kInfo];
  float4 box;
};

__device__ __forceinline__ bool bary_inside(const float (&w)[3]) {
  return w[0] <= 1 && w[0] >= 0 && w[1] <= 1 && w[1] >= 0 && w[2] <= 1 && w[2] >= 0;
}
__device__ __forceinline__ bool front_facing(const float* v) {
  return (v[7] - v[1]) * (v[3] - v[0]) < (v[4] - v[1]) * (v[6] - v[0]);
}

// Signed squared image-plane distance from the pixel to the triangle (kernel.cu:79-166): the closest point is found
// in barycentric coordinates by projecting onto an edge through the Gram matrix; inside -> sign +1, nearest of the
// three edges; outside -> sign -1, the edge picked from the signs of w (an obtuse corner redirects to the other edge).
__device__ __forceinline__ void pixel_to_face(const FaceRec& fr, const float (&w)[3], float xp, float yp, float& sign,
                                              float& dx, float& dy) {
  const float* g = fr.info + 9;
  const float* obt = fr.info + 18;
  const float* v = fr.v;
  if (w[0] > 0 && w[1] > 0 && w[2] > 0 && w[0] < 1 && w[1] < 1 && w[2] < 1) {
    float best = 100000000.f, bx = 0.f, by = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int a = k, b = (k + 1) % 3, c = (k + 2) % 3;
      const float d0 = g[3 * a + 0] - g[3 * b + 0], d1 = g[3 * a + 1] - g[3 * b + 1], d2 = g[3 * a + 2] - g[3 * b + 2];
      const float d[3] = {d0, d1, d2};
      float t[3];
      t[a] = fdiv(w[0] * d0 + w[1] * d1 + w[2] * d2 - d[b], d[a] - d[b]);
      t[b] = 1 - t[a];
      t[c] = 0;
      t[0] -= w[0]; t[1] -= w[1]; t[2] -= w[2];
      const float ex = t[0] * v[0] + t[1] * v[3] + t[2] * v[6];
      const float ey = t[0] * v[1] + t[1] * v[4] + t[2] * v[7];
      const float dd = ex * ex + ey * ey;
      if (dd < best) { best = dd; bx = ex; by = ey; }
    }
    dx = bx; dy = by; sign = 1.f;
    return;
  }
  int a = 0;  // (all w > 0 with one of them >= 1 cannot happen in exact arithmetic; the reference leaves it undefined)
  if (w[1] <= 0 && w[2] <= 0) {
    a = 0;
    if (obt[0] == 1 && (xp - v[0]) * (v[6] - v[0]) + (yp - v[1]) * (v[7] - v[1]) > 0) a = 2;
  } else if (w[2] <= 0 && w[0] <= 0) {
    a = 1;
    if (obt[1] == 1 && (xp - v[3]) * (v[0] - v[3]) + (yp - v[4]) * (v[1] - v[4]) > 0) a = 0;
  } else if (w[0] <= 0 && w[1] <= 0) {
    a = 2;
    if (obt[2] == 1 && (xp - v[6]) * (v[3] - v[6]) + (yp - v[7]) * (v[4] - v[7]) > 0) a = 1;
  } else if (w[0] <= 0) a = 1;
  else if (w[1] <= 0) a = 2;
  else if (w[2] <= 0) a = 0;
  const int b = (a + 1) % 3, c = (a + 2) % 3;
  const float d[3] = {g[3 * a + 0] - g[3 * b + 0], g[3 * a + 1] - g[3 * b + 1], g[3 * a + 2] - g[3 * b + 2]};
  float t[3];
  t[a] = fdiv(w[0] * d[0] + w[1] * d[1] + w[2] * d[2] - d[b], d[a] - d[b]);
  t[b] = 1 - t[a];
  t[c] = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) t[k] = fminf(fmaxf(t[k], 0.f), 1.f) - w[k];
  dx = t[0] * v[0] + t[1] * v[3] + t[2] * v[6];
  dy = t[0] * v[1] + t[1] * v[4] + t[2] * v[7];
  sign = -1.f;
}

__device__ __forceinline__ float sample_texture(const float* __restrict__ tex, const float (&wc)[3], int R, int T,
                                                int k, int mode) {
  if (mode == 0) {  // per-face texel grid of resolution R, folded along the diagonal (kernel.cu:196-210)
    const int wx = (int)(wc[0] * R), wy = (int)(wc[1] * R);
    int idx = ((wc[0] + wc[1]) * R - wx - wy <= 1) ? wy * R + wx : (R - 1 - wy) * R + (R - 1 - wx);
    idx = min(max(idx, 0), T - 1);  // wc[0] == 1 indexes one texel past an R x R grid in the reference
    return __ldg(tex + idx * 3 + k);
  }
  return wc[0] * __ldg(tex + k) + wc[1] * __ldg(tex + 3 + k) + wc[2] * __ldg(tex + 6 + k);
}

__global__ void __launch_bounds__(kTile* kTile) k_raster_tile(const float* __restrict__ faces,
                                                              const float* __restrict__ textures,
                                                              const float* __restrict__ info,
                                                              const float4* __restrict__ box,
                                                              const int* __restrict__ offsets,
                                                              const int* __restrict__ list, float* __restrict__ colors,
                                                              float* __restrict__ aggrs, RasterParams P) {
  __shared__ FaceRec rec[kChunk];
  __shared__ int ids[kChunk];
  const int per_img = P.tiles * P.tiles;
  const int tile = blockIdx.x, bn = tile / per_img, t = tile - bn * per_img;
  const int tx = t % P.tiles, ty = t / P.tiles;
  // a warp owns an 8 x 4 pixel footprint: a marching-cubes face (grown box ~6 x 6 pixels) then meets ~4 warps
  // instead of the ~7 of a 16 x 2 footprint, and more of a visiting warp's lanes are inside it
  constexpr int kWarpsX = kTile / kFW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int fx0 = tx * kTile + (warp % kWarpsX) * kFW, fy0 = ty * kTile + (warp / kWarpsX) * kFH;
  const int xi = fx0 + (lane % kFW), row = fy0 + (lane / kFW), is = P.is;
  const bool live = xi < is && row < is;
  const int pn = row * is + xi;
  const float xp = pix_coord(xi, is), yp = pix_coord(is - 1 - row, is);
  const long long plane = (long long)is * is;
  float* col = colors + (long long)bn * 4 * plane + pn;

  TileSpan span;  // pixel-centre extremes of the warp's footprint, clipped to the image
  {
    const int x1 = min(fx0 + kFW, is) - 1, r1 = min(fy0 + kFH, is) - 1;
    span.xlo = pix_coord(fx0, is); span.xhi = pix_coord(x1, is);
    span.ylo = pix_coord(is - 1 - r1, is); span.yhi = pix_coord(is - 1 - fy0, is);
  }
  const float threshold = P.dist_eps * P.sigma;
  const float inv_sigma = 1.f / P.sigma, inv_gamma = 1.f / P.gamma, inv_depth = 1.f / (P.far_ - P.near_);
  float c[3] = {1.f, 1.f, 1.f};
  float alpha = P.alpha == 2 ? 1.f : 0.f;
  float ssum = expf(P.eps / P.gamma), smax = P.eps;
  if (live) {
#pragma unroll
    for (int k = 0; k < 3; ++k) c[k] = col[k * plane] * (P.rgb == 1 ? ssum : 1.f);  // background
  }
  float depth_min = 10000000.f;
  int face_min = -1;

  const int beg = offsets[tile * kSeg], end = offsets[(tile + 1) * kSeg];
  const float* fbase = faces + (long long)bn * P.F * 9;
  const float* ibase = info + (long long)bn * P.F * kInfo;
  const float4* bbase = box + (long long)bn * P.F;
  const float* tbase = textures + (long long)bn * P.F * P.T * 3;
  for (int c0 = beg; c0 < end; c0 += kChunk) {
    const int n = min(kChunk, end - c0);
    __syncthreads();
    if (tid < n) ids[tid] = list[c0 + tid];
    __syncthreads();
    for (int k = tid; k < n * 9; k += kTile * kTile) rec[k / 9].v[k % 9] = __ldg(fbase + (long long)ids[k / 9] * 9 + k % 9);
    for (int k = tid; k < n * kInfo; k += kTile * kTile)
      rec[k / kInfo].info[k % kInfo] = __ldg(ibase + (long long)ids[k / kInfo] * kInfo + k % kInfo);
    if (tid < n) rec[tid].box = __ldg(bbase + ids[tid]);
    __syncthreads();
    // which of the chunk's faces can touch this warp's footprint at all (lane l tests faces l and l + 32)
    const unsigned long long wmask =
        (unsigned long long)__ballot_sync(0xffffffffu, lane < n && touches(rec[lane].box, span)) |
        ((unsigned long long)__ballot_sync(0xffffffffu, lane + 32 < n && touches(rec[(lane + 32) % kChunk].box, span))
         << 32);
    if (!live) continue;
    for (unsigned long long rest = wmask; rest; rest &= rest - 1) {
      const int j = __ffsll((long long)rest) - 1;
      const FaceRec& fr = rec[j];
      if (xp > fr.box.y || xp < fr.box.x || yp > fr.box.w || yp < fr.box.z) continue;
      float w[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) w[k] = fr.info[3 * k] * xp + fr.info[3 * k + 1] * yp + fr.info[3 * k + 2];
      float frag;
      if (P.dist == 0) {
        if (!bary_inside(w)) continue;
        frag = 1.f;
      } else if (P.dist == 1) {
        float d = w[0] > w[1] ? (w[1] > w[2] ? w[2] : w[1]) : (w[0] > w[2] ? w[2] : w[0]);
        d = d > 0 ? d * d : -(d * d);
        if (-d >= threshold) continue;
        frag = fdiv(1.f, 1.f + __expf(-d * inv_sigma));
      } else {
        float sign, dx, dy;
        pixel_to_face(fr, w, xp, yp, sign, dx, dy);
        const float d = dx * dx + dy * dy;
        if (sign < 0 && d >= threshold) continue;
        frag = fdiv(1.f, 1.f + __expf(-sign * d * inv_sigma));
      }
      if (P.alpha == 0) { if (frag > 0.5f) alpha = 1.f; }
      else if (P.alpha == 1) alpha += frag;
      else alpha *= 1.f - frag;

      float wc[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) wc[k] = fmaxf(fminf(w[k], 1.f), 0.f);
      const float wsum = fmaxf(wc[0] + wc[1] + wc[2], 1e-5f);
#pragma unroll
      for (int k = 0; k < 3; ++k) wc[k] = fdiv(wc[k], wsum);
      const float zp = fdiv(1.f, fdiv(wc[0], fr.v[2]) + fdiv(wc[1], fr.v[5]) + fdiv(wc[2], fr.v[8]));
      if (zp < P.near_ || zp > P.far_) continue;
      const float* tex = tbase + (long long)ids[j] * P.T * 3;
      if (P.rgb == 0) {
        if (zp < depth_min && bary_inside(w) && (P.double_side || front_facing(fr.v))) {
          depth_min = zp;
          face_min = ids[j];
#pragma unroll
          for (int k = 0; k < 3; ++k) c[k] = sample_texture(tex, wc, P.R, P.T, k, P.sample);
        }
      } else if (front_facing(fr.v) || P.double_side) {
        const float zn = (P.far_ - zp) * inv_depth;
        float rescale = 1.f;
        if (zn > smax) {
          rescale = __expf((smax - zn) * inv_gamma);
          smax = zn;
        }
        const float ez = __expf((zn - smax) * inv_gamma);
        ssum = rescale * ssum + ez * frag;
#pragma unroll
        for (int k = 0; k < 3; ++k)
          c[k] = rescale * c[k] + ez * frag * sample_texture(tex, wc, P.R, P.T, k, P.sample);
      }
    }
  }
  if (!live) return;
  col[3 * plane] = P.alpha == 0 ? alpha : (P.alpha == 1 ? alpha / P.F : 1.f - alpha);
  float* ag = aggrs ? aggrs + (long long)bn * 2 * plane + pn : nullptr;
  if (P.rgb == 0) {
    if (face_min != -1) {
#pragma unroll
      for (int k = 0; k < 3; ++k) col[k * plane] = c[k];
    }
    if (ag) { ag[0] = depth_min; ag[plane] = (float)face_min; }
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) col[k * plane] = c[k] / ssum;
    if (ag) { ag[0] = ssum; ag[plane] = smax; }
  }
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_soft_rasterize_tiles(int image_size) { return image_size < 1 ? -1 : (image_size + kTile - 1) / kTile; }
int lbmff_soft_rasterize_offsets(int B, int image_size) {
  const long long t = lbmff_soft_rasterize_tiles(image_size);
  const long long n = t < 1 || B < 1 ? -1 : (long long)B * t * t * kSeg + 1;
  return n > 0x7fffffffLL ? -1 : (int)n;
}

int lbmff_soft_rasterize_bin(const float* face_vertices, float* face_info, float* face_box, int* tile_offsets,
                             int B, int F, int image_size, double sigma_val, double dist_eps,
                             lbmff_stream_t stream) {
  LBMFF_REQUIRE(face_vertices && face_info && face_box && tile_offsets, "null pointer");
  LBMFF_REQUIRE(B >= 1 && F >= 1 && image_size >= 1, "non-positive size");
  LBMFF_REQUIRE((long long)B * F < 0x7fffffffLL, "too many faces");
  const int tiles = lbmff_soft_rasterize_tiles(image_size), ntile = B * tiles * tiles;
  cudaStream_t st = (cudaStream_t)stream;
  const float grow = sqrtf((float)dist_eps * (float)sigma_val);
  k_face_setup<<<(B * F + 255) / 256, 256, 0, st>>>(face_vertices, face_info, reinterpret_cast<float4*>(face_box),
                                                     B * F, grow);
  const int groups = (F + 31) / 32;
  float4* gbox = reinterpret_cast<float4*>(face_box) + (long long)B * F;  // group boxes live behind the face boxes
  k_group_box<<<(B * groups * 32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(face_box), gbox, F, groups,
                                                             B * groups);
  k_bin<false><<<(ntile * kSeg * 32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(face_box), gbox, F,
                                                                groups, image_size, tiles, ntile, tile_offsets,
                                                                nullptr);
  k_scan_exclusive<<<1, 1024, 0, st>>>(tile_offsets, ntile * kSeg);
  return check_launch(__func__);
}

int lbmff_soft_rasterize_forward(const float* face_vertices, const float* textures, const float* face_info,
                                 const float* face_box, const int* tile_offsets, int* tile_faces, float* soft_colors,
                                 float* aggrs_info, int B, int F, int image_size, int texture_size, double near_,
                                 double far_, double eps, double sigma_val, int func_id_dist, double dist_eps,
                                 double gamma_val, int func_id_rgb, int func_id_alpha, int texture_sample_type,
                                 int double_side, lbmff_stream_t stream) {
  LBMFF_REQUIRE(face_vertices && textures && face_info && face_box && tile_offsets && tile_faces && soft_colors,
                "null pointer");
  LBMFF_REQUIRE(B >= 1 && F >= 1 && image_size >= 1 && texture_size >= 1, "non-positive size");
  LBMFF_REQUIRE(func_id_dist >= 0 && func_id_dist <= 2 && func_id_rgb >= 0 && func_id_rgb <= 1 &&
                    func_id_alpha >= 0 && func_id_alpha <= 2 && texture_sample_type >= 0 && texture_sample_type <= 1,
                "unknown aggregation / distance / texture mode");
  LBMFF_REQUIRE(texture_sample_type == 0 || texture_size >= 3, "vertex textures need 3 texels per face");
  RasterParams P;
  P.B = B; P.F = F; P.is = image_size; P.T = texture_size; P.R = (int)sqrt((double)texture_size);
  if (P.R < 1) P.R = 1;
  P.near_ = (float)near_; P.far_ = (float)far_; P.eps = (float)eps; P.sigma = (float)sigma_val;
  P.dist_eps = (float)dist_eps; P.gamma = (float)gamma_val;
  P.dist = func_id_dist; P.rgb = func_id_rgb; P.alpha = func_id_alpha; P.sample = texture_sample_type;
  P.double_side = double_side ? 1 : 0;
  P.tiles = lbmff_soft_rasterize_tiles(image_size);
  const int ntile = B * P.tiles * P.tiles;
  cudaStream_t st = (cudaStream_t)stream;
  const int groups = (F + 31) / 32;
  const float4* gbox = reinterpret_cast<const float4*>(face_box) + (long long)B * F;
  k_bin<true><<<(ntile * kSeg * 32 + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(face_box), gbox, F,
                                                               groups, image_size, P.tiles, ntile,
                                                               const_cast<int*>(tile_offsets), tile_faces);
  k_raster_tile<<<ntile, kTile * kTile, 0, st>>>(face_vertices, textures, face_info,
                                                 reinterpret_cast<const float4*>(face_box), tile_offsets, tile_faces,
                                                 soft_colors, aggrs_info, P);
  return check_launch(__func__);
}

}  // extern "C"
