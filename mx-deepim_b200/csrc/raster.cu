// raster.cu -- batched triangle rasteriser (depth / RGB / mask) for sm_100a.
//
// Replaces the OpenGL renderers
//   /root/reference/lib/render_glumpy/render_py_multi.py:17-178
//   /root/reference/lib/render_glumpy/render_py_light_modelnet_multi.py:14-231
// and the post-render block
//   /root/reference/lib/pair_matching/batch_updater_py_multi.py:279-320.
//
// B200 design (one render call = 4 launches for a whole batch, any mix of meshes):
//   1. setup  : one thread per (triangle, pair): pose transform in float64, near-plane
//               clipping, pinhole projection, snap to 1/256 px, edge orientation, GL window depth;
//               writes a 64-byte record and counts the 32x32 tiles it overlaps with
//               warp-aggregated atomics (__match_any_sync on the tile id).
//   2. scan   : exclusive scan of the per-tile counts (one block per pair).
//   3. fill   : tile lists (record indices), again warp-aggregated.
//   4. raster : one CTA per (tile, pair); threads take triangles of the tile's list and
//               atomicMin a packed (bits(window depth) << 32 | record) key into a shared-memory
//               z-buffer; after one barrier every thread resolves four adjacent pixels
//               (perspective-correct attributes by ray/triangle intersection, nearest texel,
//               optional Lambert term) and writes depth, CHW RGB - mean, HWC BGR, mask with
//               128-bit stores.  Background tiles skip straight to the stores.
// The frame is written exactly once (20 B/px), the mesh is read once per pair.
//
// Compiled with -fmad=false: the float64 geometry must round exactly like the C oracle.
#include <limits.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace rc {

constexpr int RT_TILE = 32;          // tile edge in pixels
constexpr int RT_THREADS = 256;
constexpr int RT_SUB = 8;            // sub-pixel bits (GL-style 1/256 px snapping)

struct __align__(16) TriRec {  // 64 bytes = four 128-bit loads
  int x0, y0, x1, y1;    // fixed-point vertices, oriented so that the area is positive
  int x2, y2;
  int tri;               // triangle index inside its mesh | RT_FLAG_*
  int bb_min;            // pixel bbox: (y << 16) | x
  int bb_max, pad_;
  double zw0;            // GL window depth ((ndc.z + 1) / 2) at the three record vertices
  double zw1, zw2;
};

struct Renderer {
  int device, H, W, max_batch, n_meshes, max_tris, tiles_x, tiles_y, list_cap;
  float *verts, *uv, *normals;
  int32_t *tris, *mesh_tri_off, *mesh_has;  // has: bit0 uv, bit1 normals
  uint8_t *tex;
  int64_t *tex_off;
  int32_t *tex_hw;                           // [M,2]
  TriRec *recs;                              // [max_batch][2*max_tris]
  int32_t *tile_count, *tile_start, *tile_fill;  // [max_batch][n_tiles (+1)]
  int32_t *tile_list;                        // [max_batch][list_cap]
  int32_t *work_count;                       // [1] number of non-empty (tile, pair) items
  int32_t *work_list;                        // [max_batch * n_tiles] b * n_tiles + tile
};

struct Cam {
  double fx, fy, cx, cy, znear, zfar;
};

__device__ __forceinline__ long long edge_fn(long long ax, long long ay, long long bx, long long by,
                                             long long px, long long py) {
  return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}

__device__ __forceinline__ int tie_bias(long long dx, long long dy) {
  return (dy > 0 || (dy == 0 && dx < 0)) ? 0 : -1;
}

// Emit one screen-space triangle (camera-space vertices P[3]) into `rec`; returns false if culled.
constexpr int RT_FLAG_SWAPPED = 1 << 30;  // record vertices are (0, 2, 1) of the source triangle
constexpr int RT_FLAG_CLIPPED = 1 << 29;  // record is a near-clipped piece (vertices are not mesh vertices)
constexpr int RT_TRI_MASK = (1 << 29) - 1;

__device__ bool make_record(const double (*P)[3], const Cam &cam, int W, int H, int tri, bool clipped,
                            TriRec &r) {
  long long X[3], Y[3];
  double zw[3];
  // window depth of a camera-space z: ndc.z = (f+n)/(f-n) - 2fn/((f-n) z), d = (ndc.z + 1)/2
  const double pa = (cam.zfar + cam.znear) / (cam.zfar - cam.znear);
  const double pb = (2.0 * cam.zfar * cam.znear) / (cam.zfar - cam.znear);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double z = P[i][2];
    if (!(z > 0.0)) return false;
    const double u = cam.fx * P[i][0] / z + cam.cx;
    const double v = cam.fy * P[i][1] / z + cam.cy;
    if (!(fabs(u) < 1e6 && fabs(v) < 1e6)) return false;
    X[i] = (long long)rint(u * 256.0);
    Y[i] = (long long)rint(v * 256.0);
    zw[i] = ((pa * z - pb) / z + 1.0) * 0.5;
  }
  long long area = edge_fn(X[0], Y[0], X[1], Y[1], X[2], Y[2]);
  if (area == 0) return false;
  int ib = 1, ic = 2;
  if (area < 0) {
    ib = 2;
    ic = 1;
    area = -area;
  }
  const long long minx = min(X[0], min(X[1], X[2])), maxx = max(X[0], max(X[1], X[2]));
  const long long miny = min(Y[0], min(Y[1], Y[2])), maxy = max(Y[0], max(Y[1], Y[2]));
  long long px0 = (minx + 255) >> RT_SUB, px1 = maxx >> RT_SUB;
  long long py0 = (miny + 255) >> RT_SUB, py1 = maxy >> RT_SUB;
  if (px0 < 0) px0 = 0;
  if (py0 < 0) py0 = 0;
  if (px1 > W - 1) px1 = W - 1;
  if (py1 > H - 1) py1 = H - 1;
  if (px0 > px1 || py0 > py1) return false;
  r.x0 = (int)X[0]; r.y0 = (int)Y[0];
  r.x1 = (int)X[ib]; r.y1 = (int)Y[ib];
  r.x2 = (int)X[ic]; r.y2 = (int)Y[ic];
  r.zw0 = zw[0];
  r.zw1 = zw[ib];
  r.zw2 = zw[ic];
  r.pad_ = 0;
  r.tri = tri | (ib == 2 ? RT_FLAG_SWAPPED : 0) | (clipped ? RT_FLAG_CLIPPED : 0);
  r.bb_min = (int)((py0 << 16) | px0);
  r.bb_max = (int)((py1 << 16) | px1);
  return true;
}

// Warp-aggregated tile counting / slot allocation.  Every lane of the warp calls this the same
// number of times (lanes without work pass key < 0).  Returns the slot for fill (base + rank).
__device__ __forceinline__ int tile_atomic(int32_t *counters, int key, bool want_slot) {
  const unsigned lane = threadIdx.x & 31;
  const int k = key >= 0 ? key : -(int)lane - 1;  // unique dummy keys never group
  const unsigned grp = __match_any_sync(0xffffffffu, k);
  int slot = 0;
  if (key >= 0) {
    const int leader = __ffs(grp) - 1;
    int base = 0;
    if ((int)lane == leader) base = atomicAdd(counters + key, __popc(grp));
    if (want_slot) {
      base = __shfl_sync(grp, base, leader);
      slot = base + __popc(grp & ((1u << lane) - 1u));
    }
  }
  return slot;
}

// ------------------------------------------------------------------ 1. setup
// grid = (ceil(max_tris / 256), B); block = 256 (whole warps always run to the end: the tile
// loop uses full-mask warp primitives).
__global__ void __launch_bounds__(256)
raster_setup_kernel(Renderer R, const int32_t *__restrict__ cls_idx,
                    const float *__restrict__ poses, Cam cam, int n_tiles) {
  const int b = blockIdx.y;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int mesh = __ldg(cls_idx + b);
  const int t_off = __ldg(R.mesh_tri_off + mesh);
  const int n_tri = __ldg(R.mesh_tri_off + mesh + 1) - t_off;
  TriRec *recs = R.recs + (size_t)b * 2 * R.max_tris;
  int32_t *count = R.tile_count + (size_t)b * n_tiles;

  TriRec rec[2];
  int n_rec = 0;
  if (t < n_tri) {
    const float *pose = poses + b * 12;
    double Pc[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int vi = __ldg(R.tris + (size_t)(t_off + t) * 3 + i);
      const double x = (double)__ldg(R.verts + (size_t)vi * 3 + 0);
      const double y = (double)__ldg(R.verts + (size_t)vi * 3 + 1);
      const double z = (double)__ldg(R.verts + (size_t)vi * 3 + 2);
#pragma unroll
      for (int r = 0; r < 3; ++r)
        Pc[i][r] = (double)pose[r * 4 + 0] * x + (double)pose[r * 4 + 1] * y +
                   (double)pose[r * 4 + 2] * z + (double)pose[r * 4 + 3];
    }
    // near-plane clipping (Sutherland-Hodgman against z >= znear), GL primitive clipping
    double Q[4][3];
    int nq = 0;
    bool clipped_any = false;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int j = (i + 1) % 3;
      const double da = Pc[i][2] - cam.znear, db = Pc[j][2] - cam.znear;
      if (da >= 0.0) {
        Q[nq][0] = Pc[i][0]; Q[nq][1] = Pc[i][1]; Q[nq][2] = Pc[i][2];
        ++nq;
      }
      if ((da >= 0.0) != (db >= 0.0)) {
        clipped_any = true;
        const double *in = da >= 0.0 ? Pc[i] : Pc[j];
        const double *out = da >= 0.0 ? Pc[j] : Pc[i];
        const double din = da >= 0.0 ? da : db, dout = da >= 0.0 ? db : da;
        const double s = din / (din - dout);
        Q[nq][0] = in[0] + s * (out[0] - in[0]);
        Q[nq][1] = in[1] + s * (out[1] - in[1]);
        Q[nq][2] = cam.znear;
        ++nq;
      }
    }
    if (nq >= 3) {
      const double T0[3][3] = {{Q[0][0], Q[0][1], Q[0][2]}, {Q[1][0], Q[1][1], Q[1][2]},
                               {Q[2][0], Q[2][1], Q[2][2]}};
      if (make_record(T0, cam, R.W, R.H, t, nq == 4 || clipped_any, rec[n_rec])) ++n_rec;
    }
    if (nq == 4) {
      const double T1[3][3] = {{Q[0][0], Q[0][1], Q[0][2]}, {Q[2][0], Q[2][1], Q[2][2]},
                               {Q[3][0], Q[3][1], Q[3][2]}};
      if (make_record(T1, cam, R.W, R.H, t, true, rec[n_rec])) ++n_rec;
    }
  }
  // fixed slots 2t, 2t+1: deterministic layout; an unused slot is marked by bb_max < 0
  if (t < R.max_tris) {
    for (int k = 0; k < 2; ++k) {
      if (k < n_rec) recs[2 * t + k] = rec[k];
      else recs[2 * t + k].bb_max = -1;
    }
  }
  // count the tiles each record overlaps (warp-aggregated)
  for (int k = 0; k < 2; ++k) {
    int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
    if (k < n_rec) {
      tx0 = (rec[k].bb_min & 0xffff) / RT_TILE;
      ty0 = (rec[k].bb_min >> 16) / RT_TILE;
      tx1 = (rec[k].bb_max & 0xffff) / RT_TILE;
      ty1 = (rec[k].bb_max >> 16) / RT_TILE;
    }
    const int nx = tx1 - tx0 + 1, n_mine = (k < n_rec) ? nx * (ty1 - ty0 + 1) : 0;
    const int n_max = __reduce_max_sync(0xffffffffu, n_mine);
    for (int i = 0; i < n_max; ++i) {
      int key = -1;
      if (i < n_mine) key = (ty0 + i / nx) * R.tiles_x + tx0 + i % nx;
      tile_atomic(count, key, false);
    }
  }
}

// ------------------------------------------------------------------- 2. scan
// grid = B; block = 256.  Exclusive scan of tile_count -> tile_start, clears tile_fill.
__global__ void __launch_bounds__(256)
raster_scan_kernel(Renderer R, int n_tiles, int32_t *__restrict__ status) {
  const int b = blockIdx.x;
  const int32_t *count = R.tile_count + (size_t)b * n_tiles;
  int32_t *start = R.tile_start + (size_t)b * (n_tiles + 1);
  int32_t *fill = R.tile_fill + (size_t)b * n_tiles;
  __shared__ int part[256];
  const int per = (n_tiles + 255) / 256;
  const int lo = threadIdx.x * per, hi = min(lo + per, n_tiles);
  int s = 0;
  for (int i = lo; i < hi; ++i) s += count[i];
  part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int i = 0; i < 256; ++i) {
      const int v = part[i];
      part[i] = acc;
      acc += v;
    }
    start[n_tiles] = acc;
    if (acc > R.list_cap && status) atomicExch(status, 1);
  }
  __syncthreads();
  int acc = part[threadIdx.x];
  for (int i = lo; i < hi; ++i) {
    start[i] = acc;
    const int c = count[i];
    acc += c;
    fill[i] = 0;
    if (c > 0) R.work_list[atomicAdd(R.work_count, 1)] = b * n_tiles + i;  // non-empty tiles only
  }
}

// ------------------------------------------------------------------- 3. fill
// grid = (ceil(2 * max_tris / 256), B); one thread per record slot.
__global__ void __launch_bounds__(256)
raster_fill_kernel(Renderer R, const int32_t *__restrict__ cls_idx, int n_tiles) {
  const int b = blockIdx.y;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  const int mesh = __ldg(cls_idx + b);
  const int n_tri = __ldg(R.mesh_tri_off + mesh + 1) - __ldg(R.mesh_tri_off + mesh);
  const TriRec *recs = R.recs + (size_t)b * 2 * R.max_tris;
  const int32_t *start = R.tile_start + (size_t)b * (n_tiles + 1);
  int32_t *fill = R.tile_fill + (size_t)b * n_tiles;
  int32_t *list = R.tile_list + (size_t)b * R.list_cap;
  int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
  bool live = false;
  if (s < 2 * n_tri) {
    const int bb_min = recs[s].bb_min, bb_max = recs[s].bb_max;
    if (bb_max >= 0) {
      live = true;
      tx0 = (bb_min & 0xffff) / RT_TILE;
      ty0 = (bb_min >> 16) / RT_TILE;
      tx1 = (bb_max & 0xffff) / RT_TILE;
      ty1 = (bb_max >> 16) / RT_TILE;
    }
  }
  const int nx = tx1 - tx0 + 1, n_mine = live ? nx * (ty1 - ty0 + 1) : 0;
  const int n_max = __reduce_max_sync(0xffffffffu, n_mine);
  for (int i = 0; i < n_max; ++i) {
    int key = -1;
    if (i < n_mine) key = (ty0 + i / nx) * R.tiles_x + tx0 + i % nx;
    const int slot = tile_atomic(fill, key, true);
    if (key >= 0) {
      const int pos = start[key] + slot;
      if (pos < R.list_cap) list[pos] = s;
    }
  }
}

// ----------------------------------------------------------------- 4. raster
struct RasterOut {
  float *depth, *image_chw, *bgr_hwc, *mask;
  int32_t *tri_id;
  const float *light_pos, *light_int;
  float mean[3];
  float mask_thresh;
  float lin_c0, lin_c1, lin_c2;  // 2fn, f+n, f-n (render_py_multi.py:139-145, float32)
  double amb, dif;
  int mode;
};

struct Shaded {
  float c[3];  // R, G, B in 0..255 as the reference's readback delivers them
};

// Colour of pixel (x, y) on triangle `tri` of `mesh` under `pose`: perspective-correct
// attributes from the ray / triangle intersection in camera space (identical to 1/w
// interpolation of the varyings), texture2D NEAREST + CLAMP_TO_EDGE on the flipped upload
// (render_py_multi.py:75-77), optional Lambert term (render_py_light_modelnet_multi.py:50-74).
__device__ Shaded shade_pixel(const Renderer &R, const RasterOut &O, const Cam &cam, int mesh,
                              const TriRec &rec, const float *__restrict__ pose, int b, int x, int y) {
  const int tri = rec.tri & RT_TRI_MASK;
  const int t_off = __ldg(R.mesh_tri_off + mesh);
  int vi[3];
  double Pc[3][3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    vi[i] = __ldg(R.tris + (size_t)(t_off + tri) * 3 + i);
    const double px = (double)__ldg(R.verts + (size_t)vi[i] * 3 + 0);
    const double py = (double)__ldg(R.verts + (size_t)vi[i] * 3 + 1);
    const double pz = (double)__ldg(R.verts + (size_t)vi[i] * 3 + 2);
#pragma unroll
    for (int r = 0; r < 3; ++r)
      Pc[i][r] = (double)pose[r * 4 + 0] * px + (double)pose[r * 4 + 1] * py +
                 (double)pose[r * 4 + 2] * pz + (double)pose[r * 4 + 3];
  }
  double b1 = 1.0 / 3.0, b2 = 1.0 / 3.0;
  if (!(rec.tri & RT_FLAG_CLIPPED)) {
    // GL-style: screen-space barycentrics from the SNAPPED vertices (integer edge functions),
    // made perspective-correct with 1/w of the vertices.
    const long long px = (long long)x << RT_SUB, py = (long long)y << RT_SUB;
    const long long w0 = edge_fn(rec.x1, rec.y1, rec.x2, rec.y2, px, py);
    const long long w1 = edge_fn(rec.x2, rec.y2, rec.x0, rec.y0, px, py);
    const long long w2 = edge_fn(rec.x0, rec.y0, rec.x1, rec.y1, px, py);
    const double area = (double)(w0 + w1 + w2);
    const bool sw = (rec.tri & RT_FLAG_SWAPPED) != 0;
    const int ib = sw ? 2 : 1, ic = sw ? 1 : 2;
    const double q0 = ((double)w0 / area) * (1.0 / Pc[0][2]);
    const double qb = ((double)w1 / area) * (1.0 / Pc[ib][2]);
    const double qc = ((double)w2 / area) * (1.0 / Pc[ic][2]);
    const double qs = q0 + qb + qc;
    const double lb = qb / qs, lc = qc / qs;
    b1 = sw ? lc : lb;
    b2 = sw ? lb : lc;
  } else {
    // near-clipped piece: ray / source-triangle intersection (same quantity, exact vertices)
    const double d[3] = {((double)x - cam.cx) / cam.fx, ((double)y - cam.cy) / cam.fy, 1.0};
    const double e1[3] = {Pc[1][0] - Pc[0][0], Pc[1][1] - Pc[0][1], Pc[1][2] - Pc[0][2]};
    const double e2[3] = {Pc[2][0] - Pc[0][0], Pc[2][1] - Pc[0][1], Pc[2][2] - Pc[0][2]};
    const double pv[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2],
                          d[0] * e2[1] - d[1] * e2[0]};
    const double det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
    if (fabs(det) > 1e-300) {
      const double tv[3] = {-Pc[0][0], -Pc[0][1], -Pc[0][2]};
      const double qv[3] = {tv[1] * e1[2] - tv[2] * e1[1], tv[2] * e1[0] - tv[0] * e1[2],
                            tv[0] * e1[1] - tv[1] * e1[0]};
      b1 = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) / det;
      b2 = (d[0] * qv[0] + d[1] * qv[1] + d[2] * qv[2]) / det;
    }
  }
  const double b0 = 1.0 - b1 - b2;
  const int has = __ldg(R.mesh_has + mesh);
  double tu = 0.0, tvv = 0.0;
  if (has & 1) {
    tu = b0 * __ldg(R.uv + (size_t)vi[0] * 2) + b1 * __ldg(R.uv + (size_t)vi[1] * 2) +
         b2 * __ldg(R.uv + (size_t)vi[2] * 2);
    tvv = b0 * __ldg(R.uv + (size_t)vi[0] * 2 + 1) + b1 * __ldg(R.uv + (size_t)vi[1] * 2 + 1) +
          b2 * __ldg(R.uv + (size_t)vi[2] * 2 + 1);
  }
  const int Ht = __ldg(R.tex_hw + mesh * 2), Wt = __ldg(R.tex_hw + mesh * 2 + 1);
  int tx = (int)floor(tu * Wt), ty = (int)floor(tvv * Ht);
  tx = min(max(tx, 0), Wt - 1);
  ty = min(max(ty, 0), Ht - 1);
  const uint8_t *tex = R.tex + __ldg(R.tex_off + mesh) + ((size_t)(Ht - 1 - ty) * Wt + tx) * 3;
  Shaded s;
  if (O.mode == RC_RENDER_TEXTURED) {
#pragma unroll
    for (int c = 0; c < 3; ++c) s.c[c] = ((float)__ldg(tex + c) / 255.0f) * 255.0f;  // :136-137
    return s;
  }
  double n[3] = {0.0, 0.0, 0.0};
  if (has & 2) {
    double no[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      no[k] = b0 * __ldg(R.normals + (size_t)vi[0] * 3 + k) +
              b1 * __ldg(R.normals + (size_t)vi[1] * 3 + k) +
              b2 * __ldg(R.normals + (size_t)vi[2] * 3 + k);
#pragma unroll
    for (int r = 0; r < 3; ++r)
      n[r] = (double)pose[r * 4 + 0] * no[0] + (double)pose[r * 4 + 1] * no[1] +
             (double)pose[r * 4 + 2] * no[2];
  }
  // GL eye space = diag(1,-1,-1) * camera space (_get_view_mtx, render_py_multi.py:171-178)
  const double pe[3] = {b0 * Pc[0][0] + b1 * Pc[1][0] + b2 * Pc[2][0],
                        -(b0 * Pc[0][1] + b1 * Pc[1][1] + b2 * Pc[2][1]),
                        -(b0 * Pc[0][2] + b1 * Pc[1][2] + b2 * Pc[2][2])};
  const double ne[3] = {n[0], -n[1], -n[2]};
  const double l[3] = {(double)O.light_pos[b * 3 + 0] - pe[0], (double)O.light_pos[b * 3 + 1] - pe[1],
                       (double)O.light_pos[b * 3 + 2] - pe[2]};
  const double ln = sqrt(ne[0] * ne[0] + ne[1] * ne[1] + ne[2] * ne[2]);
  const double ls = sqrt(l[0] * l[0] + l[1] * l[1] + l[2] * l[2]);
  double br = (ne[0] * l[0] + ne[1] * l[1] + ne[2] * l[2]) / (ls * ln);
  br = fmax(fmin(br, 1.0), 0.0);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // ModelNet: t * ((a + b*br) * I), render_py_light_modelnet_multi.py:74;
    // LINEMOD:  t * (a + b*br * I),   render_py_light.py:74 / render_py_light_multi_program.py:76
    const double li = (double)O.light_int[b * 3 + c];
    double v = (O.mode == RC_RENDER_LIT_LINEMOD)
                   ? ((double)__ldg(tex + c) / 255.0) * (O.amb + O.dif * br * li)
                   : ((double)__ldg(tex + c) / 255.0) * ((O.amb + O.dif * br) * li);
    v = fmax(fmin(v, 1.0), 0.0);
    s.c[c] = (float)floor(v * 255.0 + 0.5);  // unorm8 colour buffer, then np.round(rgb * 255)
  }
  return s;
}

// ------------------------------------------------------------------ 4a. clear
// Streams the background of every output: depth 0, mask 0, image = 0 - mean, bgr 0, tri -1.
// Plain 128-bit stores, grid-stride; ~20 B/px at HBM write speed.
__global__ void __launch_bounds__(256)
raster_clear_kernel(RasterOut O, long long plane_quads, int n_batch) {
  const long long total = plane_quads * n_batch;  // float4 per single-channel output
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += stride) {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    st_f4(reinterpret_cast<float4 *>(O.depth) + q, z);
    if (O.mask) st_f4(reinterpret_cast<float4 *>(O.mask) + q, z);
    if (O.tri_id) reinterpret_cast<int4 *>(O.tri_id)[q] = make_int4(-1, -1, -1, -1);
    if (O.image_chw) {
      const long long b = q / plane_quads, r = q - b * plane_quads;
      float4 *img = reinterpret_cast<float4 *>(O.image_chw) + b * 3 * plane_quads + r;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float m = 0.f - O.mean[c];
        st_f4(img + c * plane_quads, make_float4(m, m, m, m));
      }
    }
    if (O.bgr_hwc) {
      float4 *o = reinterpret_cast<float4 *>(O.bgr_hwc) + q * 3;
      st_f4(o, z);
      st_f4(o + 1, z);
      st_f4(o + 2, z);
    }
  }
}

// grid = a few CTAs per SM (persistent over the non-empty tiles); block = 256; 8 KB z-buffer.
__global__ void __launch_bounds__(RT_THREADS)
raster_tile_kernel(Renderer R, const int32_t *__restrict__ cls_idx,
                   const float *__restrict__ poses, Cam cam, RasterOut O, int n_tiles, int n_batch) {
  __shared__ unsigned long long zbuf[RT_TILE * RT_TILE];
  const int W = R.W, H = R.H;
  // persistent CTAs walk the compacted list of NON-EMPTY (tile, pair) items; the background of
  // the frame is written by raster_clear_kernel at streaming speed.
  const int n_work = *R.work_count;
  for (int wi = blockIdx.x; wi < n_work; wi += gridDim.x) {
  const int work = R.work_list[wi];
  const int b = work / n_tiles, tile = work - b * n_tiles;
  const int tx0 = (tile % R.tiles_x) * RT_TILE, ty0 = (tile / R.tiles_x) * RT_TILE;
  const int32_t *start = R.tile_start + (size_t)b * (n_tiles + 1);
  const int first = start[tile];
  const int n = min(start[tile + 1], R.list_cap) - first;
  const TriRec *recs = R.recs + (size_t)b * 2 * R.max_tris;
  if (n > 0) {
    for (int i = threadIdx.x; i < RT_TILE * RT_TILE; i += RT_THREADS) zbuf[i] = ~0ull;
    __syncthreads();
    const int32_t *list = R.tile_list + (size_t)b * R.list_cap + first;
    for (int i = threadIdx.x; i < n; i += RT_THREADS) {
      const int ri = list[i];
      const int4 *rp = reinterpret_cast<const int4 *>(recs + ri);
      const int4 q0 = __ldg(rp), q1 = __ldg(rp + 1), q2 = __ldg(rp + 2), q3 = __ldg(rp + 3);
      const long long ax = q0.x, ay = q0.y, bx = q0.z, by = q0.w, cx = q1.x, cy = q1.y;
      const double zwa = __hiloint2double(q2.w, q2.z), zwb = __hiloint2double(q3.y, q3.x),
                   zwc = __hiloint2double(q3.w, q3.z);
      const int xl = max(q1.w & 0xffff, tx0), xh = min(q2.x & 0xffff, tx0 + RT_TILE - 1);
      const int yl = max(q1.w >> 16, ty0), yh = min(q2.x >> 16, ty0 + RT_TILE - 1);
      const int bias0 = tie_bias(cx - bx, cy - by), bias1 = tie_bias(ax - cx, ay - cy),
                bias2 = tie_bias(bx - ax, by - ay);
      const long long s0 = -(cy - by) * 256, s1 = -(ay - cy) * 256, s2 = -(by - ay) * 256;
      const double area = (double)edge_fn(ax, ay, bx, by, cx, cy);
      for (int y = yl; y <= yh; ++y) {
        const long long px = (long long)xl << RT_SUB, py = (long long)y << RT_SUB;
        long long w0 = edge_fn(bx, by, cx, cy, px, py);
        long long w1 = edge_fn(cx, cy, ax, ay, px, py);
        long long w2 = edge_fn(ax, ay, bx, by, px, py);
        for (int x = xl; x <= xh; ++x) {
          if (((w0 + bias0) | (w1 + bias1) | (w2 + bias2)) >= 0) {
            // window depth, interpolated linearly in screen space, stored as float32
            const double l0 = (double)w0 / area, l1 = (double)w1 / area, l2 = (double)w2 / area;
            const float d = (float)(l0 * zwa + l1 * zwb + l2 * zwc);
            if (d >= 0.0f && d <= 1.0f) {  // depth clipping of the fragment (far plane)
              const unsigned long long key =
                  ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)ri;  // GL_LESS
              atomicMin(&zbuf[(y - ty0) * RT_TILE + (x - tx0)], key);
            }
          }
          w0 += s0;
          w1 += s1;
          w2 += s2;
        }
      }
    }
    __syncthreads();
  }
  // resolve: four adjacent pixels per thread
  const int row = threadIdx.x >> 3, xq = (threadIdx.x & 7) * 4;
  const int y = ty0 + row, x = tx0 + xq;
  if (y < H && x < W) {
  const int mesh = __ldg(cls_idx + b);
  const float *pose = poses + b * 12;
  float dz[4], cr[4], cg[4], cb[4];
  int tid4[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    dz[k] = 0.f;
    cr[k] = cg[k] = cb[k] = 0.f;
    tid4[k] = -1;
    if (n > 0 && x + k < W) {
      const unsigned long long key = zbuf[row * RT_TILE + xq + k];
      if (key != ~0ull) {
        // readback linearisation, render_py_multi.py:139-145, float32 like numpy
        const float d = __uint_as_float((unsigned)(key >> 32));
        dz[k] = (d == 1.0f) ? 0.f : O.lin_c0 / (O.lin_c1 - O.lin_c2 * (2.0f * d - 1.0f));
        const TriRec rec = recs[(unsigned)key];
        tid4[k] = rec.tri & RT_TRI_MASK;
        if (O.image_chw || O.bgr_hwc) {
          const Shaded s = shade_pixel(R, O, cam, mesh, rec, pose, b, x + k, y);
          cr[k] = s.c[0];
          cg[k] = s.c[1];
          cb[k] = s.c[2];
        }
      }
    }
  }
  const size_t plane = (size_t)H * W;
  const size_t pix = (size_t)y * W + x;
  st_f4(reinterpret_cast<float4 *>(O.depth + b * plane + pix), make_float4(dz[0], dz[1], dz[2], dz[3]));
  if (O.mask) {
    const float th = O.mask_thresh;
    st_f4(reinterpret_cast<float4 *>(O.mask + b * plane + pix),
          make_float4(dz[0] > th ? 1.f : 0.f, dz[1] > th ? 1.f : 0.f, dz[2] > th ? 1.f : 0.f,
                      dz[3] > th ? 1.f : 0.f));
  }
  if (O.image_chw) {
    float *img = O.image_chw + (size_t)b * 3 * plane + pix;
    st_f4(reinterpret_cast<float4 *>(img),
          make_float4(cr[0] - O.mean[0], cr[1] - O.mean[0], cr[2] - O.mean[0], cr[3] - O.mean[0]));
    st_f4(reinterpret_cast<float4 *>(img + plane),
          make_float4(cg[0] - O.mean[1], cg[1] - O.mean[1], cg[2] - O.mean[1], cg[3] - O.mean[1]));
    st_f4(reinterpret_cast<float4 *>(img + 2 * plane),
          make_float4(cb[0] - O.mean[2], cb[1] - O.mean[2], cb[2] - O.mean[2], cb[3] - O.mean[2]));
  }
  if (O.bgr_hwc) {
    float4 *o = reinterpret_cast<float4 *>(O.bgr_hwc + ((size_t)b * plane + pix) * 3);
    st_f4(o, make_float4(cb[0], cg[0], cr[0], cb[1]));
    st_f4(o + 1, make_float4(cg[1], cr[1], cb[2], cg[2]));
    st_f4(o + 2, make_float4(cr[2], cb[3], cg[3], cr[3]));
  }
  if (O.tri_id)
    *reinterpret_cast<int4 *>(O.tri_id + b * plane + pix) = make_int4(tid4[0], tid4[1], tid4[2], tid4[3]);
  }
  if (n > 0) __syncthreads();  // z-buffer is re-initialised by the next work item
  }
}

}  // namespace rc

// ------------------------------------------------------------------ host API
namespace rc {

template <typename T>
static int dev_alloc(T **p, size_t n) {
  *p = nullptr;
  RC_CUDA_TRY(cudaMalloc(reinterpret_cast<void **>(p), (n ? n : 1) * sizeof(T)));
  return RC_OK;
}

static void renderer_free(Renderer *r) {
  if (!r) return;
  void *ptrs[] = {r->verts, r->uv, r->normals, r->tris, r->mesh_tri_off, r->mesh_has, r->tex,
                  r->tex_off, r->tex_hw, r->recs, r->work_count, r->tile_start, r->tile_fill,
                  r->tile_list, r->work_list};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  delete r;
}

}  // namespace rc

#define RC_TRY(expr)            \
  do {                          \
    int _s = (expr);            \
    if (_s != RC_OK) {          \
      rc::renderer_free(r);     \
      return rc::note_status(_s); \
    }                           \
  } while (0)

extern "C" {

int rc_renderer_create(const rc_mesh_desc *meshes, int n_meshes, int height, int width,
                       int max_batch, int device_id, void **handle) {
  RC_REQUIRE(meshes && handle && n_meshes > 0 && height > 0 && width > 0 && max_batch > 0,
             RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && width < 32768 && height < 32768, RC_ERR_UNALIGNED);
  size_t V = 0, T = 0, TX = 0;
  int max_tris = 0;
  for (int m = 0; m < n_meshes; ++m) {
    const rc_mesh_desc &d = meshes[m];
    RC_REQUIRE(d.vertices && d.triangles && d.n_vertices > 0 && d.n_triangles > 0,
               RC_ERR_INVALID_ARG);
    RC_REQUIRE(d.texture && d.tex_h > 0 && d.tex_w > 0, RC_ERR_INVALID_ARG);
    V += d.n_vertices;
    T += d.n_triangles;
    TX += (size_t)d.tex_h * d.tex_w * 3;
    if (d.n_triangles > max_tris) max_tris = d.n_triangles;
  }
  int cur = -1;
  RC_CUDA_TRY(cudaGetDevice(&cur));
  if (cur != device_id) RC_CUDA_TRY(cudaSetDevice(device_id));
  rc::Renderer *r = new rc::Renderer();
  memset(r, 0, sizeof(*r));
  r->device = device_id;
  r->H = height;
  r->W = width;
  r->max_batch = max_batch;
  r->n_meshes = n_meshes;
  r->max_tris = max_tris;
  r->tiles_x = (width + rc::RT_TILE - 1) / rc::RT_TILE;
  r->tiles_y = (height + rc::RT_TILE - 1) / rc::RT_TILE;
  const int n_tiles = r->tiles_x * r->tiles_y;
  r->list_cap = 8 * max_tris + 16 * n_tiles;
  // host staging of the concatenated mesh arrays
  std::vector<float> hv(V * 3), huv(V * 2, 0.f), hn(V * 3, 0.f);
  std::vector<int32_t> ht(T * 3), hoff(n_meshes + 1), hhas(n_meshes), hhw(n_meshes * 2);
  std::vector<int64_t> htoff(n_meshes);
  std::vector<uint8_t> htex(TX);
  size_t v0 = 0, t0 = 0, x0 = 0;
  for (int m = 0; m < n_meshes; ++m) {
    const rc_mesh_desc &d = meshes[m];
    memcpy(&hv[v0 * 3], d.vertices, sizeof(float) * 3 * d.n_vertices);
    if (d.uv) memcpy(&huv[v0 * 2], d.uv, sizeof(float) * 2 * d.n_vertices);
    if (d.normals) memcpy(&hn[v0 * 3], d.normals, sizeof(float) * 3 * d.n_vertices);
    for (int i = 0; i < d.n_triangles * 3; ++i) {
      const int idx = d.triangles[i];
      if (idx < 0 || idx >= d.n_vertices) {
        rc::renderer_free(r);
        return rc::note_status(RC_ERR_INVALID_ARG);
      }
      ht[t0 * 3 + i] = idx + (int32_t)v0;
    }
    hoff[m] = (int32_t)t0;
    hhas[m] = (d.uv ? 1 : 0) | (d.normals ? 2 : 0);
    hhw[m * 2] = d.tex_h;
    hhw[m * 2 + 1] = d.tex_w;
    htoff[m] = (int64_t)x0;
    memcpy(&htex[x0], d.texture, (size_t)d.tex_h * d.tex_w * 3);
    v0 += d.n_vertices;
    t0 += d.n_triangles;
    x0 += (size_t)d.tex_h * d.tex_w * 3;
  }
  hoff[n_meshes] = (int32_t)t0;
  RC_TRY(rc::dev_alloc(&r->verts, V * 3));
  RC_TRY(rc::dev_alloc(&r->uv, V * 2));
  RC_TRY(rc::dev_alloc(&r->normals, V * 3));
  RC_TRY(rc::dev_alloc(&r->tris, T * 3));
  RC_TRY(rc::dev_alloc(&r->mesh_tri_off, (size_t)n_meshes + 1));
  RC_TRY(rc::dev_alloc(&r->mesh_has, (size_t)n_meshes));
  RC_TRY(rc::dev_alloc(&r->tex, TX));
  RC_TRY(rc::dev_alloc(&r->tex_off, (size_t)n_meshes));
  RC_TRY(rc::dev_alloc(&r->tex_hw, (size_t)n_meshes * 2));
  RC_TRY(rc::dev_alloc(&r->recs, (size_t)max_batch * 2 * max_tris));
  RC_TRY(rc::dev_alloc(&r->work_count, (size_t)max_batch * n_tiles + 4));  // [4 | tile counts]
  r->tile_count = r->work_count + 4;
  RC_TRY(rc::dev_alloc(&r->work_list, (size_t)max_batch * n_tiles));
  RC_TRY(rc::dev_alloc(&r->tile_start, (size_t)max_batch * (n_tiles + 1)));
  RC_TRY(rc::dev_alloc(&r->tile_fill, (size_t)max_batch * n_tiles));
  RC_TRY(rc::dev_alloc(&r->tile_list, (size_t)max_batch * r->list_cap));
#define RC_UP(dst, src) \
  if (cudaMemcpy(dst, src.data(), src.size() * sizeof(src[0]), cudaMemcpyHostToDevice) != cudaSuccess) { \
    rc::renderer_free(r); \
    return rc::note_status(RC_ERR_CUDA); \
  }
  RC_UP(r->verts, hv) RC_UP(r->uv, huv) RC_UP(r->normals, hn) RC_UP(r->tris, ht)
  RC_UP(r->mesh_tri_off, hoff) RC_UP(r->mesh_has, hhas) RC_UP(r->tex, htex)
  RC_UP(r->tex_off, htoff) RC_UP(r->tex_hw, hhw)
#undef RC_UP
  *handle = r;
  return rc::note_status(RC_OK);
}

int rc_renderer_destroy(void *handle) {
  rc::renderer_free(static_cast<rc::Renderer *>(handle));
  return rc::note_status(RC_OK);
}

int rc_render_dev(void *handle, const rc_render_args *a) {
  RC_REQUIRE(handle && a, RC_ERR_INVALID_ARG);
  rc::Renderer &R = *static_cast<rc::Renderer *>(handle);
  const int B = a->batch_size;
  RC_REQUIRE(B >= 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(B <= R.max_batch, RC_ERR_CAPACITY);
  if (B == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(a->cls_idx && a->poses && a->depth, RC_ERR_INVALID_ARG);
  RC_REQUIRE(a->mode == RC_RENDER_TEXTURED || a->mode == RC_RENDER_LAMBERT ||
                 a->mode == RC_RENDER_LIT_LINEMOD,
             RC_ERR_INVALID_ARG);
  RC_REQUIRE(a->mode == RC_RENDER_TEXTURED || (a->light_pos && a->light_int), RC_ERR_INVALID_ARG);
  RC_REQUIRE(a->znear > 0.f && a->zfar > a->znear, RC_ERR_INVALID_ARG);
  RC_REQUIRE(rc::aligned16(a->depth) && rc::aligned16(a->image_chw) && rc::aligned16(a->bgr_hwc) &&
                 rc::aligned16(a->mask) && rc::aligned16(a->tri_id),
             RC_ERR_UNALIGNED);
  cudaStream_t s = static_cast<cudaStream_t>(a->stream);
  const int n_tiles = R.tiles_x * R.tiles_y;
  rc::Cam cam;
  cam.fx = a->K[0];
  cam.fy = a->K[4];
  cam.cx = a->K[2];
  cam.cy = a->K[5];
  cam.znear = a->znear;
  cam.zfar = a->zfar;
  rc::RasterOut O;
  O.depth = a->depth;
  O.image_chw = a->image_chw;
  O.bgr_hwc = a->bgr_hwc;
  O.mask = a->mask;
  O.tri_id = a->tri_id;
  O.light_pos = a->light_pos;
  O.light_int = a->light_int;
  for (int i = 0; i < 3; ++i) O.mean[i] = a->pixel_means[i];
  O.mask_thresh = a->mask_thresh;
  O.lin_c0 = (float)(2.0 * (double)a->zfar * (double)a->znear);
  O.lin_c1 = (float)((double)a->zfar + (double)a->znear);
  O.lin_c2 = (float)((double)a->zfar - (double)a->znear);
  O.amb = 1.0 - (double)a->brightness_ratio;
  O.dif = (double)a->brightness_ratio;
  O.mode = a->mode;
  auto mark = [&](int i) -> int {
    if (a->marks && a->marks[i])
      RC_CUDA_TRY(cudaEventRecord(static_cast<cudaEvent_t>(a->marks[i]), s));
    return RC_OK;
  };
  int st;
  if ((st = mark(0)) != RC_OK) return st;
  RC_CUDA_TRY(cudaMemsetAsync(R.work_count, 0, sizeof(int32_t) * ((size_t)B * n_tiles + 4), s));
  dim3 gs((R.max_tris + 255) / 256, B);
  rc::raster_setup_kernel<<<gs, 256, 0, s>>>(R, a->cls_idx, a->poses, cam, n_tiles);
  if ((st = mark(1)) != RC_OK) return st;
  rc::raster_scan_kernel<<<B, 256, 0, s>>>(R, n_tiles, a->status);
  if ((st = mark(2)) != RC_OK) return st;
  dim3 gf((2 * R.max_tris + 255) / 256, B);
  rc::raster_fill_kernel<<<gf, 256, 0, s>>>(R, a->cls_idx, n_tiles);
  if ((st = mark(3)) != RC_OK) return st;
  static int sm_count = 0;
  if (sm_count == 0 &&
      cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, R.device) != cudaSuccess)
    sm_count = 148;
  {
    const long long plane_quads = (long long)R.H * R.W / 4;
    long long blocks = (plane_quads * B + 255) / 256;
    if (blocks > sm_count * 16) blocks = sm_count * 16;
    rc::raster_clear_kernel<<<(int)blocks, 256, 0, s>>>(O, plane_quads, B);
  }
  int gr = sm_count * 4;
  if (gr > n_tiles * B) gr = n_tiles * B;
  rc::raster_tile_kernel<<<gr, rc::RT_THREADS, 0, s>>>(R, a->cls_idx, a->poses, cam, O, n_tiles, B);
  if ((st = mark(4)) != RC_OK) return st;
  return rc::note_status(rc::check_launch());
}

int rc_render_stats(void *handle, int batch_size, void *stream, int64_t *out) {
  RC_REQUIRE(handle && out && batch_size > 0, RC_ERR_INVALID_ARG);
  rc::Renderer &R = *static_cast<rc::Renderer *>(handle);
  RC_REQUIRE(batch_size <= R.max_batch, RC_ERR_CAPACITY);
  const int n_tiles = R.tiles_x * R.tiles_y;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  RC_CUDA_TRY(cudaStreamSynchronize(s));
  std::vector<int32_t> st((size_t)batch_size * (n_tiles + 1));
  RC_CUDA_TRY(cudaMemcpy(st.data(), R.tile_start, st.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
  std::vector<int32_t> bb((size_t)batch_size * 2 * R.max_tris * 16);
  RC_CUDA_TRY(cudaMemcpy(bb.data(), R.recs, bb.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
  int64_t recs = 0, entries = 0;
  for (int b = 0; b < batch_size; ++b) entries += st[(size_t)b * (n_tiles + 1) + n_tiles];
  for (size_t i = 0; i < (size_t)batch_size * 2 * R.max_tris; ++i)
    if (bb[i * 16 + 8] >= 0) ++recs;  // bb_max of live records (slots of absent triangles excluded
                                       // only when the mesh has max_tris triangles)
  out[0] = recs;
  out[1] = entries;
  return rc::note_status(RC_OK);
}

}  // extern "C"
