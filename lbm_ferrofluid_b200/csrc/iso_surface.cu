// Iso-surface extraction on the device (first half): replaces the host round trip
//     verts, faces = mcubes.marching_cubes((-phi * s).cpu().numpy()[0, 0], 0)          (src/LBM/utils/io.py:50-63)
// PyMCubes is a third-party package that is not vendored in the reference and not installed in this image, so its
// triangle table is not available; the case table used here is DERIVED at library load from first principles:
// for each of the 256 corner-sign configurations, on every cube face the cut edges are joined by directed segments
// (inside on the left seen from outside the cube; on an ambiguous face -- signs + - + - around it -- every inside
// corner is cut off on its own, a rule that depends on the face's four values only, so the two cubes sharing the
// face agree and the surface is watertight), the segments are chained into closed loops and every loop is
// fan-triangulated.  Vertices lie on cube edges at the linear zero crossing, computed from the lower-index corner so
// that the cubes sharing an edge produce the same bits.
//
// Output is a triangle soup [F][3][3] in ARRAY-INDEX coordinates (axis 0, axis 1, axis 2) = (z, y, x), the order
// PyMCubes returns; that is directly the `face_vertices` the soft rasteriser consumes.  Triangle order is
// deterministic (cell-major): two passes (count per block of one cell row -> exclusive scan -> emit), no atomics.
// Orientation: the counter-clockwise normal points towards DECREASING field values, flip = 1 reverses it.
// [PyMCubes' own convention cannot be checked here: unpinned.  The choice follows the reference's call site:
// save_rendered_image passes MINUS phi (liquid positive, gas negative, io.py:60-62) to a renderer whose only light
// shines along +y onto faces whose normal has a positive y component (lighting.py:190-196); the liquid's upper
// surface is lit only if the normals point out of the liquid, i.e. towards decreasing values of the volume passed.]
#include "capi.cuh"

#include <mutex>

namespace lbmff {

constexpr int kMcMaxTri = 8;  // the generator asserts no case needs more (the classic table needs 5)

struct McTable {
  signed char ntri[256];
  signed char edge[256][kMcMaxTri * 3];  // cube edge ids, 3 per triangle
};

// corner c = bit0: +1 along axis 2 (x), bit1: +1 along axis 1 (y), bit2: +1 along axis 0 (z)
// edge e = axis * 4 + (b0 | b1 << 1): runs along `axis` (0 = x, 1 = y, 2 = z), (b0, b1) = offsets along the other two
// axes in increasing axis order.
static void edge_corners(int e, int& c0, int& c1) {
  const int a = e / 4, b0 = e & 1, b1 = (e >> 1) & 1;
  const int o0 = a == 0 ? 1 : 0, o1 = a == 2 ? 1 : 2;  // the two other axes, increasing
  c0 = (b0 << o0) | (b1 << o1);
  c1 = c0 | (1 << a);
}
static int edge_between(int ca, int cb) {
  for (int e = 0; e < 12; ++e) {
    int c0, c1;
    edge_corners(e, c0, c1);
    if ((c0 == ca && c1 == cb) || (c0 == cb && c1 == ca)) return e;
  }
  return -1;
}

// do the three cube edges lie on one face of the cube?
static bool edges_on_one_face(int e0, int e1, int e2) {
  int c[6];
  edge_corners(e0, c[0], c[1]);
  edge_corners(e1, c[2], c[3]);
  edge_corners(e2, c[4], c[5]);
  for (int ax = 0; ax < 3; ++ax)
    for (int side = 0; side < 2; ++side) {
      bool all = true;
      for (int k = 0; k < 6; ++k) all = all && (((c[k] >> ax) & 1) == side);
      if (all) return true;
    }
  return false;
}

static bool build_mc_table(McTable& T) {
  // the six faces: corners in counter-clockwise order seen from OUTSIDE the cube (right-handed x, y, z = bits 0, 1, 2)
  int faces[6][4];
  int nf = 0;
  for (int a = 0; a < 3; ++a)
    for (int side = 0; side < 2; ++side) {
      const int u = (a + 1) % 3, v = (a + 2) % 3;  // (a, u, v) is a cyclic permutation: u x v = +a
      const int base = side << a;
      const int ccw[4] = {base, base | (1 << u), base | (1 << u) | (1 << v), base | (1 << v)};  // CCW around +a
      for (int k = 0; k < 4; ++k) faces[nf][k] = side ? ccw[k] : ccw[(4 - k) % 4];               // outward = -a: reverse
      ++nf;
    }
  for (int cfg = 0; cfg < 256; ++cfg) {
    int next[12];
    for (int e = 0; e < 12; ++e) next[e] = -1;
    for (int f = 0; f < 6; ++f) {
      const int* c = faces[f];
      for (int i = 0; i < 4; ++i) {
        const int si = (cfg >> c[i]) & 1, sn = (cfg >> c[(i + 1) % 4]) & 1;
        if (!(si == 1 && sn == 0)) continue;  // an inside -> outside crossing, walking counter-clockwise
        int j = i;
        int guard = 0;
        while (((cfg >> c[(j + 3) % 4]) & 1) && guard++ < 4) j = (j + 3) % 4;  // back over the inside corners
        const int e_out = edge_between(c[i], c[(i + 1) % 4]);
        const int e_in = edge_between(c[(j + 3) % 4], c[j]);
        if (e_out < 0 || e_in < 0 || next[e_out] != -1) return false;
        next[e_out] = e_in;  // inside lies to the left of e_out -> e_in seen from outside the cube
      }
    }
    int nt = 0;
    bool used[12] = {false};
    for (int e0 = 0; e0 < 12; ++e0) {
      if (next[e0] < 0 || used[e0]) continue;
      int loop[12], n = 0;
      for (int e = e0; !used[e]; e = next[e]) {
        if (next[e] < 0 || n >= 12) return false;  // an open chain: the construction is broken
        used[e] = true;
        loop[n++] = e;
      }
      // triangulate the loop so that no triangle lies IN a cube face (three cut edges of one face: an ambiguous face
      // visited twice by the loop) -- such a flat triangle would overlap its mirror image from the neighbouring cube.
      // Candidates: the fan from every apex, then the zigzag strip from every start.
      int best[3 * 10], nbest = -1;
      for (int variant = 0; variant < 2 && nbest < 0; ++variant)
        for (int rot = 0; rot < n && nbest < 0; ++rot) {
          int t[3 * 10], m = 0;
          auto L = [&](int k) { return loop[(k + rot) % n]; };
          if (variant == 0) {
            for (int k = 1; k + 1 < n; ++k) { t[m++] = L(0); t[m++] = L(k); t[m++] = L(k + 1); }
          } else {
            int lo = 0, hi = n - 1;
            bool take_lo = true;
            while (hi - lo >= 2) {
              if (take_lo) { t[m++] = L(lo); t[m++] = L(lo + 1); t[m++] = L(hi); ++lo; }
              else { t[m++] = L(lo); t[m++] = L(hi - 1); t[m++] = L(hi); --hi; }
              take_lo = !take_lo;
            }
          }
          bool flat = false;
          for (int k = 0; k < m; k += 3) flat = flat || edges_on_one_face(t[k], t[k + 1], t[k + 2]);
          if (!flat) { nbest = m; for (int k = 0; k < m; ++k) best[k] = t[k]; }
        }
      if (nbest < 0) return false;
      for (int k = 0; k < nbest; k += 3) {
        if (nt >= kMcMaxTri) return false;
        T.edge[cfg][3 * nt + 0] = (signed char)best[k];
        T.edge[cfg][3 * nt + 1] = (signed char)best[k + 1];
        T.edge[cfg][3 * nt + 2] = (signed char)best[k + 2];
        ++nt;
      }
    }
    T.ntri[cfg] = (signed char)nt;
    for (int k = 3 * nt; k < kMcMaxTri * 3; ++k) T.edge[cfg][k] = -1;
  }
  return true;
}

static McTable g_host_table;
static bool g_table_ok = false;
static std::once_flag g_table_once;
static const McTable* host_table() {
  std::call_once(g_table_once, [] { g_table_ok = build_mc_table(g_host_table); });
  return g_table_ok ? &g_host_table : nullptr;
}

__device__ McTable d_mc_table;
static std::mutex g_upload_mutex;
static bool g_uploaded[64] = {false};
static cudaError_t ensure_device_table() {
  const McTable* T = host_table();
  if (!T) return cudaErrorUnknown;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(g_upload_mutex);
  if (dev >= 0 && dev < 64 && g_uploaded[dev]) return cudaSuccess;
  e = cudaMemcpyToSymbol(d_mc_table, T, sizeof(McTable));
  if (e == cudaSuccess && dev >= 0 && dev < 64) g_uploaded[dev] = true;
  return e;
}

// ---- device side -------------------------------------------------------------------------------------------
struct McGrid {
  int D, H, W;  // field extents; cells are (D-1) x (H-1) x (W-1)
};

__device__ __forceinline__ int mc_config(const float* __restrict__ u, const McGrid g, int x, int y, int z, float iso,
                                         float (&val)[8]) {
  int cfg = 0;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int xx = x + (c & 1), yy = y + ((c >> 1) & 1), zz = z + (c >> 2);
    val[c] = __ldg(u + ((long long)zz * g.H + yy) * g.W + xx);
    if (val[c] < iso) cfg |= 1 << c;  // "inside" = below the iso value
  }
  return cfg;
}

// Blocks: blockIdx.x = cell row (z * (H-1) + y), blockIdx.y = 256-cell chunk of the row; the block's rank in
// cell-major order is row * chunks + chunk, so triangle order stays cell-major, and a thread's cell costs one
// 32-bit division per block instead of two 64-bit ones per thread.
struct McCell {
  int x, y, z, rank;
  bool live;
};
__device__ __forceinline__ McCell mc_cell(const McGrid g) {
  McCell c;
  const int ch = g.H - 1, row = blockIdx.x;
  c.z = row / ch;
  c.y = row - c.z * ch;
  c.x = blockIdx.y * 256 + threadIdx.x;
  c.live = c.x < g.W - 1;
  c.rank = row * gridDim.y + blockIdx.y;
  return c;
}

// pass 1: triangles per block
__global__ void __launch_bounds__(256) k_mc_count(const float* __restrict__ u, McGrid g, float iso,
                                                  int* __restrict__ block_count) {
  const McCell c = mc_cell(g);
  int n = 0;
  if (c.live) {
    float val[8];
    n = d_mc_table.ntri[mc_config(u, g, c.x, c.y, c.z, iso, val)];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
  __shared__ int sm[8];
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += sm[w];
    block_count[c.rank] = t;
  }
}

// exclusive scan of the block counts by ONE block (in place), total -> *total
__global__ void __launch_bounds__(1024) k_mc_scan(int* __restrict__ counts, int n, int* __restrict__ total) {
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
      int w = warp_tot[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_tot[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int before = carry + (warp ? warp_tot[warp - 1] : 0) + (s - v);
    if (i < n) counts[i] = before;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

__device__ __forceinline__ void mc_edge_vertex(int e, int x, int y, int z, const float (&val)[8], float iso,
                                               float (&p)[3]) {
  const int a = e >> 2, b0 = e & 1, b1 = (e >> 1) & 1;
  const int o0 = a == 0 ? 1 : 0, o1 = a == 2 ? 1 : 2;
  const int c0 = (b0 << o0) | (b1 << o1), c1 = c0 | (1 << a);
  const float v0 = val[c0], v1 = val[c1];
  float t = (iso - v0) / (v1 - v0);  // c0 is the lower-index corner in every cube that shares this edge
  t = fminf(fmaxf(t, 0.0f), 1.0f);
  const float px = (float)(x + (c0 & 1)) + (a == 0 ? t : 0.0f);
  const float py = (float)(y + ((c0 >> 1) & 1)) + (a == 1 ? t : 0.0f);
  const float pz = (float)(z + (c0 >> 2)) + (a == 2 ? t : 0.0f);
  p[0] = pz; p[1] = py; p[2] = px;  // array-index order (axis 0, 1, 2), as PyMCubes returns
}

// pass 2: emit; tri [F][3][3]
__global__ void __launch_bounds__(256) k_mc_emit(const float* __restrict__ u, McGrid g, float iso, int flip,
                                                 const int* __restrict__ block_offset, float* __restrict__ tri,
                                                 int capacity) {
  const McCell c = mc_cell(g);
  const int x = c.x, y = c.y, z = c.z;
  int n = 0, cfg = 0;
  float val[8];
  if (c.live) {
    cfg = mc_config(u, g, x, y, z, iso, val);
    n = d_mc_table.ntri[cfg];
  }
  // exclusive scan of n inside the block
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int s = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, s, o);
    if (lane >= o) s += t;
  }
  __shared__ int sm[8];
  if (lane == 31) sm[warp] = s;
  __syncthreads();
  int before = block_offset[c.rank] + (s - n);
  for (int w = 0; w < warp; ++w) before += sm[w];
  for (int t = 0; t < n; ++t) {
    if (before + t >= capacity) return;
    float* out = tri + (long long)(before + t) * 9;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      // the table's loops run with "inside" (below iso) on the left seen from outside the cube: their right-hand
      // normal points towards increasing values (tests/test_iso_surface.py); the default orientation is the reverse
      const int kk = flip ? k : (k == 0 ? 0 : 3 - k);
      float p[3];
      mc_edge_vertex(d_mc_table.edge[cfg][3 * t + kk], x, y, z, val, iso, p);
      out[3 * k + 0] = p[0]; out[3 * k + 1] = p[1]; out[3 * k + 2] = p[2];
    }
  }
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_mc_case_table(int* ntri, int* edges) {
  const McTable* T = host_table();
  if (!T) return bad_arg(__func__, "case-table construction failed");
  if (!(ntri && edges)) return bad_arg(__func__, "null pointer");
  for (int c = 0; c < 256; ++c) {
    ntri[c] = T->ntri[c];
    for (int k = 0; k < kMcMaxTri * 3; ++k) edges[c * kMcMaxTri * 3 + k] = T->edge[c][k];
  }
  return 0;
}

int lbmff_iso_surface_blocks(int D, int H, int W) {
  if (D < 2 || H < 2 || W < 2) return -1;
  const long long nb = (long long)(D - 1) * (H - 1) * ((W - 1 + 255) / 256);
  return nb > 0x7fffffffLL ? -1 : (int)nb;
}

int lbmff_iso_surface_count_3d(const float* u, double iso, int* block_offsets, int* total, int D, int H, int W,
                               lbmff_stream_t stream) {
  LBMFF_REQUIRE(u && block_offsets && total, "null pointer");
  const int nb = lbmff_iso_surface_blocks(D, H, W);
  LBMFF_REQUIRE(nb > 0, "every extent must be >= 2 (and the cell count below 2^39)");
  cudaError_t e = ensure_device_table();
  if (e != cudaSuccess) { set_last_error(__func__, e); return (int)e; }
  const McGrid g{D, H, W};
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid((unsigned)((D - 1) * (H - 1)), (unsigned)((W - 1 + 255) / 256));
  k_mc_count<<<grid, 256, 0, st>>>(u, g, (float)iso, block_offsets);
  k_mc_scan<<<1, 1024, 0, st>>>(block_offsets, nb, total);
  return check_launch(__func__);
}

int lbmff_iso_surface_emit_3d(const float* u, double iso, int flip, const int* block_offsets, float* triangles,
                              int capacity, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(u && block_offsets && (triangles || capacity == 0), "null pointer");
  LBMFF_REQUIRE(capacity >= 0, "negative capacity");
  const int nb = lbmff_iso_surface_blocks(D, H, W);
  LBMFF_REQUIRE(nb > 0, "every extent must be >= 2 (and the cell count below 2^39)");
  if (capacity == 0) return 0;
  cudaError_t e = ensure_device_table();
  if (e != cudaSuccess) { set_last_error(__func__, e); return (int)e; }
  const McGrid g{D, H, W};
  const dim3 grid((unsigned)((D - 1) * (H - 1)), (unsigned)((W - 1 + 255) / 256));
  k_mc_emit<<<grid, 256, 0, (cudaStream_t)stream>>>(u, g, (float)iso, flip, block_offsets, triangles, capacity);
  return check_launch(__func__);
}

}  // extern "C"
