// raster.cu -- batched triangle rasteriser (depth / RGB / mask) for sm_100a.
//
// Replaces the OpenGL renderers
//   /root/reference/lib/render_glumpy/render_py_multi.py:17-178
//   /root/reference/lib/render_glumpy/render_py_light_modelnet_multi.py:14-231
// and the post-render block
//   /root/reference/lib/pair_matching/batch_updater_py_multi.py:279-320.
//
// B200 design (one render call = 7 launches for a whole batch, any mix of meshes, no memset;
// launch order: clear, vertex, setup, scan, fill, cover, resolve):
//   clear   : the background (depth 0, mask 0, image 0 - mean ...) in row order with 128-bit
//             stores -- the whole frame, or in incremental mode (a caller that owns its frame
//             buffers) only the tiles the previous frame covered.  Launched first; it releases
//             its dependents at once and the passes below run beside it (programmatic dependent
//             launches, each waiting for its predecessor only); only resolve waits for it.
//   vertex  : one thread per (vertex, pair): pose transform in float64, pinhole projection, snap
//             to 1/256 px, GL window depth; also zeroes this call's counters.
//   setup   : one thread per (triangle, pair): edge orientation, culling (incl. micro-triangles
//             that contain no pixel centre).  A MICRO-TRIANGLE (pixel box <= 3x3: most of a dense
//             mesh) that does contain pixel centres is finished here: its fragments go straight
//             into the pair's dense z-winner blocks with a 64-bit atomicMin (RED in L2) on the
//             packed (bits(window depth) << 32 | slot) key and its tile gets a mark -- no record,
//             no tile-list entry, no coverage work.  Larger triangles get a 64-byte record + a
//             packed tile box, and the 16x16 tiles they overlap are counted with warp-aggregated
//             atomics (__match_any_sync on the tile id).  Triangles straddling the near plane go
//             to a list that the scan blocks clip (usually empty).
//   scan    : near-plane clipping of the pair's straddlers, exclusive scan of the per-tile counts
//             (one block per pair), the list of non-empty (tile, pair) items (list or mark) and
//             the list of items that have a tile list.
//   fill    : tile lists (record indices), again warp-aggregated, one thread per triangle.
//   cover   : per item with a tile list, atomicMin of the packed key into a shared-memory
//             z-buffer; large triangles are tested cooperatively, one pixel per thread; the
//             winners are merged into the dense z-winner blocks.
//   resolve : one thread per pixel of the non-empty tiles: reads the winner and puts the empty
//             key back, perspective-correct attributes from the projected vertices, nearest
//             texel, optional Lambert term, fused BGR->RGB / CHW / -mean and mask.
// Every pixel of the frame ends up written by exactly one of clear / resolve; the mesh is read
// once per pair.
//
// Compiled with -fmad=false: no compiler-chosen contraction in the float64 geometry (the results
// are checked against the C oracle's literal GL pipeline at the snapping / float32 level).
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace rc {

constexpr int RT_TW = 16;            // tile width  in pixels (multiple of 4, power of two)
constexpr int RT_TH = 16;            // tile height in pixels
constexpr int RT_PIX = RT_TW * RT_TH;
constexpr int RT_TW_SHIFT = 4;       // log2(RT_TW)
constexpr int RT_THREADS = 256;
constexpr int RT_SUB = 8;            // sub-pixel bits (GL-style 1/256 px snapping)
static_assert((1 << RT_TW_SHIFT) == RT_TW && RT_TW % 4 == 0, "tile width");
static_assert(RT_THREADS % (RT_PIX / 4) == 0, "whole tiles per background pass");
static_assert(RT_TH == RT_TW, "one shift for tile rows and columns");

struct __align__(16) TriRec {  // 64 bytes = four 128-bit loads
  int x0, y0, x1, y1;    // fixed-point vertices, oriented so that the area is positive
  int x2, y2;
  int tri;               // triangle index inside its mesh | RT_FLAG_*
  int bb_min;            // pixel bbox: (y << 16) | x
  int bb_max, pad_;
  double zw0;            // GL window depth ((ndc.z + 1) / 2) at the three record vertices
  double zw1, zw2;
};

// One projected vertex: snapped screen position (1/256 px), GL window depth.  In the vertex
// buffer the first word carries the flags in its two low bits: (X << 2) | RT_VF_*.
struct __align__(16) VtxS {
  int X, Y;
  double zw;
};
constexpr int RT_VF_VALID = 1;  // in front of the camera and |u|, |v| < 1e6 px
constexpr int RT_VF_NEAR = 2;   // z >= znear (inside the near plane)
constexpr unsigned RT_TBOX_NONE = 0xffu;  // packed tile box with tx0 = 255 > tx1 = 0: no record
// tx0 = 254 > tx1 = 0: a micro-triangle that went straight to the z-winner blocks (no record, no
// tile-list entry); only rc_render_stats tells it from RT_TBOX_NONE
constexpr unsigned RT_TBOX_DIRECT = 0xfeu;

struct Renderer {
  int device, H, W, max_batch, n_meshes, max_tris, tiles_x, tiles_y, list_cap;
  float *verts, *uv, *normals;
  int32_t *tris, *mesh_tri_off, *mesh_has;  // has: bit0 uv, bit1 normals
  int32_t *mesh_vert_off;                    // [M+1] first vertex of each mesh in the concatenated arrays
  int max_verts;
  VtxS *vtx;                                 // [max_batch][max_verts] projected vertices of this call
  double *viz;                               // [max_batch][max_verts] their 1 / camera-space z (perspective-correct attributes)
  uint8_t *tex;
  int64_t *tex_off;
  int32_t *tex_hw;                           // [M,2]
  TriRec *recs;                              // [max_batch][2*max_tris]
  uint32_t *tbox;                            // [max_batch][max_tris] tile box of slot 2t: tx0 | ty0<<8 | tx1<<16 | ty1<<24
  int32_t *clip_list;                        // [max_batch * max_tris] b * max_tris + t of near-plane straddlers
  uint2 *clip_box;                           // [max_batch * max_tris] tile boxes of the two pieces of clip_list[i]
  int32_t *tile_count, *tile_start, *tile_fill;  // [max_batch][n_tiles (+1)]
  int32_t *tile_touch;                       // [B][n_tiles] this call's "a micro-triangle wrote here" marks (behind tile_count)
  int32_t *bin_list;                         // [max_batch * n_tiles] the items that also have a tile list (coverage pass)
  int32_t *bin_tri, *bin_n;                  // [max_batch][max_tris], [max_batch]: the pair's triangles that got a record (fill pass)
  int32_t *tile_list;                        // [max_batch][list_cap]
  int32_t *work_count;                       // [RT_WC_INTS] header (see RT_WC_*), then tile_count
  unsigned long long *zwin;                  // [max_batch][n_tiles][RT_PIX] winning keys; all ~0 between frames (resolve resets what it reads)
  int32_t *work_list;                        // [max_batch * n_tiles] packed non-empty (tile, pair) items (this call's half)
  int32_t *heavy;                            // [1] their number (this call's half)
  const int32_t *prev_list, *prev_heavy;     // the previous call's items = tiles its frame dirtied
  // ---- host-side state (not used by kernels): ping-pong work lists for the incremental background
  int32_t *work_list_b[2], *heavy_count;     // heavy_count [2]
  int cur, have_prev, last_B, n_frames;
  const void *last_out[5];
  float last_mean[3];
};

// work_count header
constexpr int RT_WC_CLIP = 3;      // clip-list length (setup)
constexpr int RT_WC_TICKET_H = 4;  // next binned item (coverage)
constexpr int RT_WC_BINNED = 5;    // number of items with a tile list (scan)
constexpr int RT_WC_INTS = 8;

// work item: pair << 16 | tile row << 8 | tile column
__device__ __forceinline__ int pack_work(int b, int ty, int tx) { return (b << 16) | (ty << 8) | tx; }

struct Cam {
  double fx, fy, cx, cy, znear, zfar;
};

// Edge function on snapped coordinates.  |X|, |Y| < 2^28 (|u|, |v| < 1e6 px at 1/256 px), so every
// difference fits in 32 bits and each product is ONE widening multiply (IMAD.WIDE) instead of a
// 64 x 64-bit multiply sequence.
__device__ __forceinline__ long long edge_fn(int ax, int ay, int bx, int by, int px, int py) {
  return (long long)(bx - ax) * (long long)(py - ay) - (long long)(by - ay) * (long long)(px - ax);
}

__device__ __forceinline__ int tie_bias(int dx, int dy) {
  return (dy > 0 || (dy == 0 && dx < 0)) ? 0 : -1;
}

constexpr int RT_FLAG_SWAPPED = 1 << 30;  // record vertices are (0, 2, 1) of the source triangle
constexpr int RT_FLAG_CLIPPED = 1 << 29;  // record is a near-clipped piece (vertices are not mesh vertices)
constexpr int RT_TRI_MASK = (1 << 29) - 1;

struct RecQ {  // a TriRec as the four 128-bit words it is stored and loaded as
  int4 q0, q1, q2, q3;
  int bb_min, bb_max;  // bb_max < 0: nothing to draw
};

// Camera-space point -> VtxS.  One reciprocal per vertex (the perspective divide as hardware does
// it); FP64 runs at half rate on sm_100 and a division is ~30 instructions, so they are counted.
__device__ __forceinline__ int project_vertex(double x, double y, double z, const Cam &cam, VtxS &o,
                                              double *iz_out = nullptr) {
  if (iz_out) *iz_out = 0.0;
  o.X = 0;
  o.Y = 0;
  o.zw = 0.0;
  int flag = (z - cam.znear >= 0.0) ? RT_VF_NEAR : 0;
  if (!(z > 0.0)) return flag;
  // window depth of a camera-space z: ndc.z = (f+n)/(f-n) - 2fn/((f-n) z), d = (ndc.z + 1)/2
  const double pa = (cam.zfar + cam.znear) / (cam.zfar - cam.znear);
  const double pb = (2.0 * cam.zfar * cam.znear) / (cam.zfar - cam.znear);
  const double iz = 1.0 / z;
  if (iz_out) *iz_out = iz;
  const double u = cam.fx * x * iz + cam.cx, v = cam.fy * y * iz + cam.cy;
  if (!(fabs(u) < 1e6 && fabs(v) < 1e6)) return flag;
  o.X = __double2int_rn(u * 256.0);  // < 2^28 in magnitude
  o.Y = __double2int_rn(v * 256.0);
  o.zw = ((pa - pb * iz) + 1.0) * 0.5;
  return flag | RT_VF_VALID;
}

// Where the micro-triangles of one pair go: straight into the pair's dense z-winner blocks.
struct DirectZ {
  unsigned long long *zwin;  // [n_tiles][RT_PIX] of this pair
  int32_t *touch;            // [n_tiles] marks of this pair
  int tiles_x;
};

constexpr int RT_REC_NONE = 0;    // nothing to draw
constexpr int RT_REC_BINNED = 1;  // r holds a record for the tile lists
constexpr int RT_REC_DIRECT = 2;  // fragments already merged into the z-winner blocks

// Screen-space record of one triangle from its three projected vertices.  Culls what cannot
// produce a fragment: degenerate, off-screen, or -- for boxes of at most 3x3 pixels -- no pixel
// centre inside under the very rule the coverage kernel applies (integer edge functions + top-left
// tie bias).  With `dz` the fragments of such a micro-triangle (the bulk of a dense mesh: a
// 20k-triangle model covers ~6k pixels) are depth-tested at once with a 64-bit atomicMin on the
// same (bits(window depth) << 32 | slot) key the coverage kernel uses, so they need no record,
// no tile-list entry and no coverage pass: the key order decides the winner, whoever writes it.
// Everything stays in registers: no dynamically indexed arrays.
__device__ __forceinline__ int make_record(const VtxS &va, const VtxS &vb, const VtxS &vc, int W,
                                           int H, int tri, bool clipped, RecQ &r,
                                           const DirectZ *dz = nullptr) {
  r.bb_max = -1;
  r.bb_min = 0;
  const int X0 = va.X, Y0 = va.Y;
  int X1 = vb.X, Y1 = vb.Y, X2 = vc.X, Y2 = vc.Y;
  const double zw0 = va.zw;
  double zw1 = vb.zw, zw2 = vc.zw;
  const long long area = edge_fn(X0, Y0, X1, Y1, X2, Y2);
  if (area == 0) return RT_REC_NONE;
  const bool sw = area < 0;
  if (sw) {  // orient so that the area is positive: record vertices (0, 2, 1)
    int t = X1; X1 = X2; X2 = t;
    t = Y1; Y1 = Y2; Y2 = t;
    const double z = zw1; zw1 = zw2; zw2 = z;
  }
  const int minx = min(X0, min(X1, X2)), maxx = max(X0, max(X1, X2));
  const int miny = min(Y0, min(Y1, Y2)), maxy = max(Y0, max(Y1, Y2));
  int px0 = (minx + 255) >> RT_SUB, px1 = maxx >> RT_SUB;
  int py0 = (miny + 255) >> RT_SUB, py1 = maxy >> RT_SUB;
  if (px0 < 0) px0 = 0;
  if (py0 < 0) py0 = 0;
  if (px1 > W - 1) px1 = W - 1;
  if (py1 > H - 1) py1 = H - 1;
  if (px0 > px1 || py0 > py1) return RT_REC_NONE;
  if (px1 - px0 <= 2 && py1 - py0 <= 2 && maxx - minx < 1024 && maxy - miny < 1024) {
    // micro-triangle: test the (at most nine) pixel centres now.  Relative to the box corner every
    // coordinate is below 2^11 in magnitude, so the edge functions fit 32 bits (same values as
    // the 64-bit ones the coverage pass evaluates: the test is exact, not conservative).
    const int qx = px0 << RT_SUB, qy = py0 << RT_SUB;
    const int x0 = X0 - qx, y0 = Y0 - qy, x1 = X1 - qx, y1 = Y1 - qy, x2 = X2 - qx, y2 = Y2 - qy;
    const int dx0 = x2 - x1, dy0 = y2 - y1, dx1 = x0 - x2, dy1 = y0 - y2, dx2 = x1 - x0,
              dy2 = y1 - y0;
    const int t1 = tie_bias(dx1, dy1), t2 = tie_bias(dx2, dy2);
    // edge functions at the corner (0, 0), then stepped: +256 in x adds -dy * 256, in y dx * 256
    const int a0i = dy0 * x1 - dx0 * y1 + tie_bias(dx0, dy0);
    const int a1i = dy1 * x2 - dx1 * y2 + t1;
    const int a2i = dy2 * x0 - dx2 * y0 + t2;
    unsigned cov = 0;  // bit 3 j + i: pixel centre (px0 + i, py0 + j) is inside
    const int ny = py1 - py0, nx = px1 - px0;
    int a0 = a0i, a1 = a1i, a2 = a2i;  // edge values at the start of the pixel row
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      int b0 = a0, b1 = a1, b2 = a2;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        if (j <= ny && i <= nx && ((b0 | b1 | b2) >= 0)) cov |= 1u << (3 * j + i);
        b0 -= dy0 << RT_SUB;
        b1 -= dy1 << RT_SUB;
        b2 -= dy2 << RT_SUB;
      }
      a0 += dx0 << RT_SUB;
      a1 += dx1 << RT_SUB;
      a2 += dx2 << RT_SUB;
    }
    if (!cov) return RT_REC_NONE;
    if (dz) {
      // window depth exactly as the coverage pass forms it: one division per triangle, then
      // d = float(zwa + w1 * dzb + w2 * dzc) with the unbiased edge values (this file is
      // compiled with -fmad=false, so the association is the written one in both places)
      const double inv_area = 1.0 / (double)(sw ? -area : area);
      const double dzb = (zw1 - zw0) * inv_area, dzc = (zw2 - zw0) * inv_area;
      const int w1c = a1i - t1, w2c = a2i - t2;  // w1, w2 at the box corner
      const unsigned slot = 2u * (unsigned)tri;
      while (cov) {
        const int bit = __ffs(cov) - 1;
        cov &= cov - 1;
        const int j = bit / 3, i = bit - 3 * j;
        const int w1 = w1c + j * (dx1 << RT_SUB) - i * (dy1 << RT_SUB);
        const int w2 = w2c + j * (dx2 << RT_SUB) - i * (dy2 << RT_SUB);
        const float d = (float)(zw0 + (double)w1 * dzb + (double)w2 * dzc);
        if (d >= 0.0f && d <= 1.0f) {  // depth clipping of the fragment (far plane)
          const int x = px0 + i, y = py0 + j;
          const int tile = (y >> RT_TW_SHIFT) * dz->tiles_x + (x >> RT_TW_SHIFT);
          const unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | slot;
          atomicMin(dz->zwin + (size_t)tile * RT_PIX + ((y & (RT_TH - 1)) << RT_TW_SHIFT) +
                        (x & (RT_TW - 1)),
                    key);
          dz->touch[tile] = 1;
        }
      }
      return RT_REC_DIRECT;
    }
  }
  r.bb_min = (int)((py0 << 16) | px0);
  r.bb_max = (int)((py1 << 16) | px1);
  r.q0 = make_int4(X0, Y0, X1, Y1);
  r.q1 = make_int4(X2, Y2,
                   tri | (sw ? RT_FLAG_SWAPPED : 0) | (clipped ? RT_FLAG_CLIPPED : 0), r.bb_min);
  r.q2 = make_int4(r.bb_max, 0, __double2loint(zw0), __double2hiint(zw0));
  r.q3 = make_int4(__double2loint(zw1), __double2hiint(zw1), __double2loint(zw2),
                   __double2hiint(zw2));
  return RT_REC_BINNED;
}

__device__ __forceinline__ void store_record(TriRec *dst, const RecQ &r) {
  int4 *p = reinterpret_cast<int4 *>(dst);
  p[0] = r.q0;
  p[1] = r.q1;
  p[2] = r.q2;
  p[3] = r.q3;
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

// Packed tile box of a pixel box ((y << 16) | x corners; bb_max < 0: none).
__device__ __forceinline__ unsigned tile_box(int bb_min, int bb_max) {
  if (bb_max < 0) return RT_TBOX_NONE;
  return (unsigned)((bb_min & 0xffff) / RT_TW) | ((unsigned)((bb_min >> 16) / RT_TH) << 8) |
         ((unsigned)((bb_max & 0xffff) / RT_TW) << 16) | ((unsigned)((bb_max >> 16) / RT_TH) << 24);
}

// The tiles of a packed tile box, visited in lock step by the whole warp.
template <typename F>
__device__ __forceinline__ void for_each_tile(unsigned tb, int tiles_x, F &&f) {
  const int tx0 = tb & 255, ty0 = (tb >> 8) & 255, tx1 = (tb >> 16) & 255, ty1 = tb >> 24;
  const int nx = tx1 - tx0 + 1;  // <= 0 for RT_TBOX_NONE
  const int n_mine = nx > 0 ? nx * (ty1 - ty0 + 1) : 0;
  const int n_max = __reduce_max_sync(0xffffffffu, n_mine);
  for (int i = 0; i < n_max; ++i) {
    int key = -1;
    if (i < n_mine) key = (ty0 + i / nx) * tiles_x + tx0 + i % nx;
    f(key);
  }
}

// ------------------------------------------------------------------ 1. setup
// Near-plane clipping (Sutherland-Hodgman against z >= znear, GL primitive clipping) of a
// triangle that straddles the plane: rare, kept out of line so that its dynamically indexed
// arrays do not cost the common path a stack frame.
__device__ __forceinline__ uint2 clip_and_emit(const double *Pc /* [3][3] */, const Cam &cam, int W,
                                           int H, int t, TriRec *slots /* 2t, 2t+1 */) {
  double Q[4][3];
  int nq = 0;
  bool clipped_any = false;
  for (int i = 0; i < 3; ++i) {
    const int j = (i + 1) % 3;
    const double *Pi = Pc + 3 * i, *Pj = Pc + 3 * j;
    const double da = Pi[2] - cam.znear, db = Pj[2] - cam.znear;
    if (da >= 0.0) {
      Q[nq][0] = Pi[0]; Q[nq][1] = Pi[1]; Q[nq][2] = Pi[2];
      ++nq;
    }
    if ((da >= 0.0) != (db >= 0.0)) {
      clipped_any = true;
      const double *in = da >= 0.0 ? Pi : Pj;
      const double *out = da >= 0.0 ? Pj : Pi;
      const double din = da >= 0.0 ? da : db, dout = da >= 0.0 ? db : da;
      const double s = din / (din - dout);
      Q[nq][0] = in[0] + s * (out[0] - in[0]);
      Q[nq][1] = in[1] + s * (out[1] - in[1]);
      Q[nq][2] = cam.znear;
      ++nq;
    }
  }
  // live records occupy the slots in order (slot 2t first)
  uint2 bb = make_uint2(RT_TBOX_NONE, RT_TBOX_NONE);
  int n_rec = 0;
  for (int k = 0; k < 2; ++k) {
    if (k == 0 ? nq < 3 : nq != 4) continue;
    const int i1 = k == 0 ? 1 : 2, i2 = k == 0 ? 2 : 3;
    RecQ r;
    r.bb_max = -1;
    VtxS pv[3];
    const int f0 = project_vertex(Q[0][0], Q[0][1], Q[0][2], cam, pv[0]);
    const int f1 = project_vertex(Q[i1][0], Q[i1][1], Q[i1][2], cam, pv[1]);
    const int f2 = project_vertex(Q[i2][0], Q[i2][1], Q[i2][2], cam, pv[2]);
    if (f0 & f1 & f2 & RT_VF_VALID)
      make_record(pv[0], pv[1], pv[2], W, H, t, k == 1 || nq == 4 || clipped_any, r);
    if (r.bb_max < 0) continue;
    store_record(slots + n_rec, r);
    if (n_rec == 0)
      bb.x = tile_box(r.bb_min, r.bb_max);
    else
      bb.y = tile_box(r.bb_min, r.bb_max);
    ++n_rec;
  }
  return bb;
}

// 0. vertices: one thread per (vertex, pair): pose transform in float64, projection, snapping,
// window depth -- once per vertex instead of once per triangle corner (a closed mesh has ~6
// corners per vertex).  grid = (ceil(max_verts / 256), B); block = 256.
__global__ void __launch_bounds__(256)
raster_vertex_kernel(Renderer R, const int32_t *__restrict__ cls_idx,
                     const float *__restrict__ poses, Cam cam, int n_tiles) {
  pdl_launch_dependents();  // setup waits (griddepcontrol.wait) for this grid's end
  const int b = blockIdx.y;
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  // this call's counters start from zero (was two memset nodes in front of the render): the
  // pair's tile counts and marks, the header and this half's item count.  Their last readers
  // (the previous frame's consumers) finished before the background pass in front of this
  // kernel could start, and their first writer (setup) waits for this whole grid.
  for (int i = v; i < n_tiles; i += gridDim.x * blockDim.x) {
    R.tile_count[(size_t)b * n_tiles + i] = 0;
    R.tile_touch[(size_t)b * n_tiles + i] = 0;
  }
  if (b == 0 && v < RT_WC_INTS) R.work_count[v] = 0;
  if (b == 0 && v == RT_WC_INTS) *R.heavy = 0;
  if (v == 0) R.bin_n[b] = 0;
  const int mesh = __ldg(cls_idx + b);
  const int v_off = __ldg(R.mesh_vert_off + mesh);
  if (v >= __ldg(R.mesh_vert_off + mesh + 1) - v_off) return;
  const float *pose = poses + b * 12;
  const double x = (double)__ldg(R.verts + (size_t)(v_off + v) * 3 + 0);
  const double y = (double)__ldg(R.verts + (size_t)(v_off + v) * 3 + 1);
  const double z = (double)__ldg(R.verts + (size_t)(v_off + v) * 3 + 2);
  double Pc[3];
#pragma unroll
  for (int r = 0; r < 3; ++r)
    Pc[r] = (double)__ldg(pose + r * 4 + 0) * x + (double)__ldg(pose + r * 4 + 1) * y +
            (double)__ldg(pose + r * 4 + 2) * z + (double)__ldg(pose + r * 4 + 3);
  VtxS o;
  double iz;
  const int flag = project_vertex(Pc[0], Pc[1], Pc[2], cam, o, &iz);
  *reinterpret_cast<int4 *>(R.vtx + (size_t)b * R.max_verts + v) =
      make_int4((o.X << 2) | flag, o.Y, __double2loint(o.zw), __double2hiint(o.zw));
  R.viz[(size_t)b * R.max_verts + v] = iz;  // resolve: 1/w of the varyings, the very same double
}

// 1. setup.  grid = (ceil(max_tris / (256 * RT_SETUP_PER)), B); block = 256 (whole warps always
// run to the end: the tile loop uses full-mask warp primitives).  One thread per RT_SETUP_PER
// (triangle, pair)s, 256 apart: the kernel is bound by the latency of its dependent gathers
// (indices -> projected vertices), so each thread keeps several in flight.  Integer work only:
// culling, the 64-byte record as four 128-bit stores, the packed tile box (what the fill pass
// re-reads: 4 bytes per triangle), and the 16x16 tiles counted with warp-aggregated atomics.
constexpr int RT_SETUP_PER = 2;

__global__ void __launch_bounds__(256, 4)
raster_setup_kernel(Renderer R, const int32_t *__restrict__ cls_idx, int n_tiles) {
  pdl_wait();               // the projected vertices
  pdl_launch_dependents();  // the clip kernel waits for this grid's end in turn
  const int b = blockIdx.y;
  const int t_base = blockIdx.x * (256 * RT_SETUP_PER) + threadIdx.x;
  const int mesh = __ldg(cls_idx + b);
  const int t_off = __ldg(R.mesh_tri_off + mesh);
  const int n_tri = __ldg(R.mesh_tri_off + mesh + 1) - t_off;
  const int v_off = __ldg(R.mesh_vert_off + mesh);
  TriRec *recs = R.recs + (size_t)b * 2 * R.max_tris;
  uint32_t *tbox = R.tbox + (size_t)b * R.max_tris;
  int32_t *count = R.tile_count + (size_t)b * n_tiles;
  const VtxS *vtx = R.vtx + (size_t)b * R.max_verts - v_off;  // indexed by the global vertex id
  DirectZ dz;
  dz.zwin = R.zwin + (size_t)b * n_tiles * RT_PIX;
  dz.touch = R.tile_touch + (size_t)b * n_tiles;
  dz.tiles_x = R.tiles_x;

  int idx[RT_SETUP_PER][3];
#pragma unroll
  for (int k = 0; k < RT_SETUP_PER; ++k) {
    const int t = t_base + k * 256;
#pragma unroll
    for (int c = 0; c < 3; ++c)
      idx[k][c] = t < n_tri ? __ldg(R.tris + (size_t)(t_off + t) * 3 + c) : v_off;
  }
  int4 pv[RT_SETUP_PER][3];
#pragma unroll
  for (int k = 0; k < RT_SETUP_PER; ++k)
#pragma unroll
    for (int c = 0; c < 3; ++c) pv[k][c] = *reinterpret_cast<const int4 *>(vtx + idx[k][c]);

  unsigned tb[RT_SETUP_PER];
#pragma unroll
  for (int k = 0; k < RT_SETUP_PER; ++k) {
    const int t = t_base + k * 256;
    tb[k] = RT_TBOX_NONE;  // slot 2t unused
    if (t < n_tri) {
      const int f_and = pv[k][0].x & pv[k][1].x & pv[k][2].x;
      const int f_or = pv[k][0].x | pv[k][1].x | pv[k][2].x;
      if (f_and & RT_VF_NEAR) {
        if (f_and & RT_VF_VALID) {
          VtxS va, vb, vc;
          va.X = pv[k][0].x >> 2; va.Y = pv[k][0].y; va.zw = __hiloint2double(pv[k][0].w, pv[k][0].z);
          vb.X = pv[k][1].x >> 2; vb.Y = pv[k][1].y; vb.zw = __hiloint2double(pv[k][1].w, pv[k][1].z);
          vc.X = pv[k][2].x >> 2; vc.Y = pv[k][2].y; vc.zw = __hiloint2double(pv[k][2].w, pv[k][2].z);
          RecQ r;
          const int kind = make_record(va, vb, vc, R.W, R.H, t, false, r, &dz);
          if (kind == RT_REC_BINNED) {
            store_record(recs + 2 * t, r);
            tb[k] = tile_box(r.bb_min, r.bb_max);
          } else if (kind == RT_REC_DIRECT) {
            tb[k] = RT_TBOX_DIRECT;
          }
        }
      } else if (f_or & RT_VF_NEAR) {
        // straddles the near plane (rare): left to raster_clip_kernel, which needs dynamically
        // indexed polygons and would cost this kernel its registers
        const int slot = atomicAdd(R.work_count + RT_WC_CLIP, 1);
        R.clip_list[slot] = b * R.max_tris + t;
      }
    }
    if (t < R.max_tris) tbox[t] = tb[k];
    // the few triangles with a record (a dense mesh: ~3 %) go on the pair's list for the fill
    // pass, one atomic per warp
    const bool binned = tb[k] != RT_TBOX_NONE && tb[k] != RT_TBOX_DIRECT;
    const unsigned bal = __ballot_sync(0xffffffffu, binned);
    if (bal) {
      const int lane = threadIdx.x & 31, leader = __ffs(bal) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(R.bin_n + b, __popc(bal));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (binned) R.bin_tri[(size_t)b * R.max_tris + base + __popc(bal & ((1u << lane) - 1u))] = t;
    }
  }
  // count the tiles each record overlaps (warp-aggregated)
#pragma unroll
  for (int k = 0; k < RT_SETUP_PER; ++k)
    for_each_tile(tb[k], R.tiles_x, [&](int key) { tile_atomic(count, key, false); });
}

// The triangles raster_setup_kernel found straddling the near plane (usually none), done by the
// pair's scan block in front of its scan (was a launch of its own: one more dependent kernel on
// the chain of every render for a list that is almost always empty).  Every block walks the
// whole clip list and keeps the entries of its own pair: clip, emit up to two records into the
// triangle's slots, count their tiles.
__device__ __noinline__ void clip_pair(const Renderer &R, const int32_t *__restrict__ cls_idx,
                                       const float *__restrict__ poses, const Cam &cam, int n_tiles,
                                       int b, int n) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int item = R.clip_list[i];
    if (item / R.max_tris != b) continue;
    const int t = item - b * R.max_tris;
    const int t_off = __ldg(R.mesh_tri_off + __ldg(cls_idx + b));
    const float *pose = poses + b * 12;
    double Pc[9];
    for (int k = 0; k < 3; ++k) {
      const int vi = __ldg(R.tris + (size_t)(t_off + t) * 3 + k);
      const double x = (double)__ldg(R.verts + (size_t)vi * 3 + 0);
      const double y = (double)__ldg(R.verts + (size_t)vi * 3 + 1);
      const double z = (double)__ldg(R.verts + (size_t)vi * 3 + 2);
      for (int r = 0; r < 3; ++r)
        Pc[k * 3 + r] = (double)__ldg(pose + r * 4 + 0) * x + (double)__ldg(pose + r * 4 + 1) * y +
                        (double)__ldg(pose + r * 4 + 2) * z + (double)__ldg(pose + r * 4 + 3);
    }
    const uint2 bb = clip_and_emit(Pc, cam, R.W, R.H, t, R.recs + (size_t)b * 2 * R.max_tris + 2 * t);
    R.clip_box[i] = bb;  // setup left tbox[t] = none: the fill pass places these from the clip list
    int32_t *count = R.tile_count + (size_t)b * n_tiles;
    for (int k = 0; k < 2; ++k) {
      const unsigned q = k ? bb.y : bb.x;
      const int tx0 = q & 255, ty0 = (q >> 8) & 255, tx1 = (q >> 16) & 255, ty1 = q >> 24;
      for (int ty = ty0; ty <= ty1; ++ty)
        for (int tx = tx0; tx <= tx1; ++tx)  // empty for RT_TBOX_NONE
          atomicAdd(count + ty * R.tiles_x + tx, 1);
    }
  }
}

// ------------------------------------------------------------------- 2. scan
__device__ __forceinline__ int block_excl_scan(int v, int *smem /* [8] */, int &total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  __syncthreads();  // smem free (a previous scan's readers are done)
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const int s = smem[w];
    if (w < warp) base += s;
    tot += s;
  }
  total = tot;
  return base + inc - v;
}

// grid = B; block = 256.  Near-plane clipping of the pair's straddlers (if any), then the
// exclusive scan of tile_count -> tile_start; clears tile_fill; appends the pair's non-empty
// tiles -- a tile list OR a micro-triangle mark -- to the work list (resolve, the next frame's
// incremental background, the compare step's sparse passes) and those with a tile list to the
// coverage pass's list (one atomic per pair and list; the order inside a pair is the tile
// order).  Leaves tile_count > 0 exactly for the tiles on the work list.
__global__ void __launch_bounds__(256)
raster_scan_kernel(Renderer R, const int32_t *__restrict__ cls_idx,
                   const float *__restrict__ poses, Cam cam, int n_tiles,
                   int32_t *__restrict__ status) {
  pdl_wait();
  pdl_launch_dependents();
  const int b = blockIdx.x;
  const int n_clip = R.work_count[RT_WC_CLIP];
  if (n_clip > 0) {  // block-uniform
    clip_pair(R, cls_idx, poses, cam, n_tiles, b, n_clip);
    __syncthreads();
  }
  int32_t *count = R.tile_count + (size_t)b * n_tiles;
  const int32_t *touch = R.tile_touch + (size_t)b * n_tiles;
  int32_t *start = R.tile_start + (size_t)b * (n_tiles + 1);
  int32_t *fill = R.tile_fill + (size_t)b * n_tiles;
  __shared__ int part[8];
  __shared__ int s_base, s_base_bin;
  const int per = (n_tiles + 255) / 256;
  const int lo = min((int)threadIdx.x * per, n_tiles), hi = min(lo + per, n_tiles);
  int s = 0, ne = 0, nb = 0;
  for (int i = lo; i < hi; ++i) {
    const int c = count[i];
    s += c;
    nb += c > 0;
    ne += (c > 0) | (touch[i] != 0);
  }
  int total, total_ne, total_nb;
  int acc = block_excl_scan(s, part, total);
  const int ne_off = block_excl_scan(ne, part, total_ne);
  const int nb_off = block_excl_scan(nb, part, total_nb);
  if (threadIdx.x == 0) {
    start[n_tiles] = total;
    if (total > R.list_cap && status) atomicExch(status, 1);
    s_base = atomicAdd(R.heavy, total_ne);
    s_base_bin = total_nb ? atomicAdd(R.work_count + RT_WC_BINNED, total_nb) : 0;
  }
  __syncthreads();
  int front = s_base + ne_off, front_bin = s_base_bin + nb_off;
  for (int i = lo; i < hi; ++i) {
    start[i] = acc;
    const int c = count[i];
    acc += c;
    fill[i] = 0;
    const int item = pack_work(b, i / R.tiles_x, i % R.tiles_x);
    if (c > 0) {
      R.bin_list[front_bin++] = item;
      R.work_list[front++] = item;
    } else if (touch[i] != 0) {
      R.work_list[front++] = item;
      count[i] = 1;  // "on the work list" for the sparse passes of the compare step
    }
  }
}

// ------------------------------------------------------------------- 3. fill
// grid = (a few, B); block = 256, striding over the pair's list of triangles that have a record
// (written by setup): reads the triangle's packed tile box (4 bytes) and places record slot 2t;
// the pieces of near-clipped triangles (slots 2t, 2t + 1 of the few triangles on the clip list)
// are placed from the clip list by the first threads of the grid.
__global__ void __launch_bounds__(256)
raster_fill_kernel(Renderer R, const int32_t *__restrict__ cls_idx, int n_tiles) {
  pdl_wait();
  pdl_launch_dependents();
  const int b = blockIdx.y;
  const int32_t *start = R.tile_start + (size_t)b * (n_tiles + 1);
  int32_t *fill = R.tile_fill + (size_t)b * n_tiles;
  int32_t *list = R.tile_list + (size_t)b * R.list_cap;
  const int n_bin = R.bin_n[b];
  const int32_t *bin = R.bin_tri + (size_t)b * R.max_tris;
  for (int base = blockIdx.x * blockDim.x; base < n_bin; base += gridDim.x * blockDim.x) {
    const int i = base + threadIdx.x;
    int t = 0;
    unsigned tb = RT_TBOX_NONE;
    if (i < n_bin) {
      t = bin[i];
      tb = R.tbox[(size_t)b * R.max_tris + t];
    }
    // Each lane walks its own tile box, four tiles per round: the four slot atomics (and the
    // four loads of the tiles' starts) are in flight together.  The warp-aggregated variant
    // (tile_atomic) needs each atomic's result for a shuffle before the next tile can be
    // started -- with ~360 listed triangles per pair there is no contention to aggregate away,
    // only a chain of L2 round trips (17 us at 128 pairs for a few microseconds of work).
    const int tx0 = tb & 255, ty0 = (tb >> 8) & 255, tx1 = (tb >> 16) & 255, ty1 = tb >> 24;
    const int nx = tx1 - tx0 + 1;  // <= 0 for RT_TBOX_NONE
    const int n_mine = nx > 0 ? nx * (ty1 - ty0 + 1) : 0;
    for (int i0 = 0; i0 < n_mine; i0 += 4) {
      int key[4], slot[4], first[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = i0 + j;
        key[j] = i < n_mine ? (ty0 + i / nx) * R.tiles_x + tx0 + i % nx : -1;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        slot[j] = key[j] >= 0 ? atomicAdd(fill + key[j], 1) : 0;
        first[j] = key[j] >= 0 ? start[key[j]] : 0;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (key[j] >= 0 && first[j] + slot[j] < R.list_cap) list[first[j] + slot[j]] = 2 * t;
    }
  }
  const int n_clip = R.work_count[RT_WC_CLIP];
  const int gid = (blockIdx.y * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid < n_clip) {
    const int stride = gridDim.x * gridDim.y * blockDim.x;
    for (int i = gid; i < n_clip; i += stride) {
      const int item = R.clip_list[i];
      const int cb = item / R.max_tris, ct = item - cb * R.max_tris;
      const uint2 bb = R.clip_box[i];
      const int32_t *cstart = R.tile_start + (size_t)cb * (n_tiles + 1);
      int32_t *cfill = R.tile_fill + (size_t)cb * n_tiles;
      int32_t *clist = R.tile_list + (size_t)cb * R.list_cap;
      for (int k = 0; k < 2; ++k) {
        const unsigned q = k ? bb.y : bb.x;
        const int tx0 = q & 255, ty0 = (q >> 8) & 255, tx1 = (q >> 16) & 255, ty1 = q >> 24;
        for (int ty = ty0; ty <= ty1; ++ty)
          for (int tx = tx0; tx <= tx1; ++tx) {
            const int key = ty * R.tiles_x + tx;
            const int pos = cstart[key] + atomicAdd(cfill + key, 1);
            if (pos < R.list_cap) clist[pos] = 2 * ct + k;
          }
      }
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

// The winning triangle of a pixel as resolve needs it: oriented snapped vertices (record order),
// flags, mesh vertex ids.  Rebuilt from the projected vertices of this call for every triangle
// that was not near-clipped (micro-triangles have no record at all); clipped pieces come from
// their record.
struct WinTri {
  int x0, y0, x1, y1, x2, y2;
  int tri;    // triangle index | RT_FLAG_*
  int vi[3];  // global vertex ids of the SOURCE triangle
  double iz[3];  // 1 / camera-space z of the source vertices (valid unless RT_FLAG_CLIPPED)
};

__device__ __forceinline__ void load_winner(const Renderer &R, int b, int mesh, unsigned slot,
                                            WinTri &w) {
  const int t = (int)(slot >> 1);
  const int t_off = __ldg(R.mesh_tri_off + mesh), v_off = __ldg(R.mesh_vert_off + mesh);
#pragma unroll
  for (int i = 0; i < 3; ++i) w.vi[i] = __ldg(R.tris + (size_t)(t_off + t) * 3 + i);
  const VtxS *vtx = R.vtx + (size_t)b * R.max_verts - v_off;
  const int4 p0 = *reinterpret_cast<const int4 *>(vtx + w.vi[0]);
  const int4 p1 = *reinterpret_cast<const int4 *>(vtx + w.vi[1]);
  const int4 p2 = *reinterpret_cast<const int4 *>(vtx + w.vi[2]);
  if ((p0.x & p1.x & p2.x & (RT_VF_NEAR | RT_VF_VALID)) == (RT_VF_NEAR | RT_VF_VALID)) {
    w.x0 = p0.x >> 2; w.y0 = p0.y;
    w.x1 = p1.x >> 2; w.y1 = p1.y;
    w.x2 = p2.x >> 2; w.y2 = p2.y;
    w.tri = t;
    const double *viz = R.viz + (size_t)b * R.max_verts - v_off;
#pragma unroll
    for (int i = 0; i < 3; ++i) w.iz[i] = viz[w.vi[i]];
    if (edge_fn(w.x0, w.y0, w.x1, w.y1, w.x2, w.y2) < 0) {  // make_record's orientation
      int q = w.x1; w.x1 = w.x2; w.x2 = q;
      q = w.y1; w.y1 = w.y2; w.y2 = q;
      w.tri |= RT_FLAG_SWAPPED;
    }
  } else {  // a piece of a near-clipped triangle: its vertices are not mesh vertices
    const int4 *rp = reinterpret_cast<const int4 *>(R.recs + (size_t)b * 2 * R.max_tris + slot);
    const int4 q0 = __ldg(rp), q1 = __ldg(rp + 1);
    w.x0 = q0.x; w.y0 = q0.y; w.x1 = q0.z; w.y1 = q0.w;
    w.x2 = q1.x; w.y2 = q1.y;
    w.tri = q1.z;
    w.iz[0] = w.iz[1] = w.iz[2] = 0.0;
  }
}

// Colour of pixel (x, y) on the winning triangle under `pose`: perspective-correct
// attributes (1/w interpolation of the varyings), texture2D NEAREST + CLAMP_TO_EDGE on the
// flipped upload (render_py_multi.py:75-77), optional Lambert term
// (render_py_light_modelnet_multi.py:50-74).  LIT = false is the textured renderer: only the
// camera-space z of the three vertices is needed (9 float64 products instead of 27).
template <bool LIT>
__device__ __forceinline__ Shaded shade_pixel(const Renderer &R, const RasterOut &O, const Cam &cam,
                                              int mesh, const WinTri &w,
                                              const float *__restrict__ pose, int b, int x, int y) {
  double Pc[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  const bool clipped = (w.tri & RT_FLAG_CLIPPED) != 0;
  // camera-space vertices: only the lit renderers and near-clipped pieces need them; a textured
  // pixel gets by with the vertex pass's 1/z (w.iz: the same double 1.0 / Pc[i][2] would give)
  if (LIT || clipped) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double px = (double)__ldg(R.verts + (size_t)w.vi[i] * 3 + 0);
      const double py = (double)__ldg(R.verts + (size_t)w.vi[i] * 3 + 1);
      const double pz = (double)__ldg(R.verts + (size_t)w.vi[i] * 3 + 2);
#pragma unroll
      for (int r = 0; r < 3; ++r)
        Pc[i][r] = (double)pose[r * 4 + 0] * px + (double)pose[r * 4 + 1] * py +
                   (double)pose[r * 4 + 2] * pz + (double)pose[r * 4 + 3];
    }
  }
  double b1 = 1.0 / 3.0, b2 = 1.0 / 3.0;
  if (!clipped) {
    // GL-style: screen-space barycentrics from the SNAPPED vertices (integer edge functions),
    // made perspective-correct with 1/w of the vertices.
    const int px = x << RT_SUB, py = y << RT_SUB;
    const long long w0 = edge_fn(w.x1, w.y1, w.x2, w.y2, px, py);
    const long long w1 = edge_fn(w.x2, w.y2, w.x0, w.y0, px, py);
    const long long w2 = edge_fn(w.x0, w.y0, w.x1, w.y1, px, py);
    const bool sw = (w.tri & RT_FLAG_SWAPPED) != 0;
    // q_i = l_i / z_i up to the common factor 1 / area, which cancels in q_i / sum(q)
    const double q0 = (double)w0 * w.iz[0];
    const double qb = (double)w1 * (sw ? w.iz[2] : w.iz[1]);
    const double qc = (double)w2 * (sw ? w.iz[1] : w.iz[2]);
    const double iqs = 1.0 / (q0 + qb + qc);
    const double lb = qb * iqs, lc = qc * iqs;
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
    tu = b0 * __ldg(R.uv + (size_t)w.vi[0] * 2) + b1 * __ldg(R.uv + (size_t)w.vi[1] * 2) +
         b2 * __ldg(R.uv + (size_t)w.vi[2] * 2);
    tvv = b0 * __ldg(R.uv + (size_t)w.vi[0] * 2 + 1) + b1 * __ldg(R.uv + (size_t)w.vi[1] * 2 + 1) +
          b2 * __ldg(R.uv + (size_t)w.vi[2] * 2 + 1);
  }
  const int Ht = __ldg(R.tex_hw + mesh * 2), Wt = __ldg(R.tex_hw + mesh * 2 + 1);
  int tx = (int)floor(tu * Wt), ty = (int)floor(tvv * Ht);
  tx = min(max(tx, 0), Wt - 1);
  ty = min(max(ty, 0), Ht - 1);
  const uint8_t *tex = R.tex + __ldg(R.tex_off + mesh) + ((size_t)(Ht - 1 - ty) * Wt + tx) * 3;
  Shaded s;
  if (!LIT) {
#pragma unroll
    for (int c = 0; c < 3; ++c) s.c[c] = ((float)__ldg(tex + c) / 255.0f) * 255.0f;  // :136-137
    return s;
  }
  double n[3] = {0.0, 0.0, 0.0};
  if (has & 2) {
    double no[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      no[k] = b0 * __ldg(R.normals + (size_t)w.vi[0] * 3 + k) +
              b1 * __ldg(R.normals + (size_t)w.vi[1] * 3 + k) +
              b2 * __ldg(R.normals + (size_t)w.vi[2] * 3 + k);
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

// ------------------------------------------------- 4. background, coverage, resolve
// Three launches write the frame exactly once:
//   4a clear   : streams the background (depth 0, mask 0, image 0 - mean, bgr 0, tri -1) in row
//                order with 128-bit stores, skipping the quads of non-empty tiles.  Bandwidth
//                bound; it releases its dependents at once and runs on a few CTAs per SM, so
//   4b cover   : (launched with programmatic stream serialisation, resident beside 4a) works
//                through the non-empty (tile, pair) items: threads take triangles of the tile's
//                list and atomicMin a packed (bits(window depth) << 32 | record) key into a
//                shared-memory z-buffer; triangles with a large pixel box inside the tile are
//                queued and tested by every thread at its own pixels instead.  Latency bound.
//                The winning keys of the tile go to a compact buffer (2 KB per item).
//   4c resolve : one thread per pixel of the non-empty tiles, no barriers: depth linearisation,
//                shading, fused image post-processing, stores.
constexpr int RT_CV_THREADS = 128;
constexpr int RT_BIG_AREA = 32;   // pixel-box area (inside the tile) above which a triangle is rasterised by the whole CTA
constexpr int RT_BIG_CAP = 192;   // queue of such triangles per tile (overflow: the owning thread loops)
static_assert(RT_PIX % RT_CV_THREADS == 0 && RT_PIX == RT_THREADS, "pixels per thread");

__device__ __forceinline__ void store_quad(const RasterOut &O, int b, int H, int W, int y, int x,
                                           const float4 dz, const float4 cr, const float4 cg,
                                           const float4 cb, const int4 tid4) {
  const size_t plane = (size_t)H * W;
  const size_t pix = (size_t)y * W + x;
  st_f4(reinterpret_cast<float4 *>(O.depth + b * plane + pix), dz);
  if (O.mask) {
    const float th = O.mask_thresh;
    st_f4(reinterpret_cast<float4 *>(O.mask + b * plane + pix),
          make_float4(dz.x > th ? 1.f : 0.f, dz.y > th ? 1.f : 0.f, dz.z > th ? 1.f : 0.f,
                      dz.w > th ? 1.f : 0.f));
  }
  if (O.image_chw) {
    float *img = O.image_chw + (size_t)b * 3 * plane + pix;
    st_f4(reinterpret_cast<float4 *>(img),
          make_float4(cr.x - O.mean[0], cr.y - O.mean[0], cr.z - O.mean[0], cr.w - O.mean[0]));
    st_f4(reinterpret_cast<float4 *>(img + plane),
          make_float4(cg.x - O.mean[1], cg.y - O.mean[1], cg.z - O.mean[1], cg.w - O.mean[1]));
    st_f4(reinterpret_cast<float4 *>(img + 2 * plane),
          make_float4(cb.x - O.mean[2], cb.y - O.mean[2], cb.z - O.mean[2], cb.w - O.mean[2]));
  }
  if (O.bgr_hwc) {
    float4 *o = reinterpret_cast<float4 *>(O.bgr_hwc + ((size_t)b * plane + pix) * 3);
    st_f4(o, make_float4(cb.x, cg.x, cr.x, cb.y));
    st_f4(o + 1, make_float4(cg.y, cr.y, cb.z, cg.z));
    st_f4(o + 2, make_float4(cr.z, cb.w, cg.w, cr.w));
  }
  if (O.tri_id) *reinterpret_cast<int4 *>(O.tri_id + b * plane + pix) = tid4;
}

// 4a.  One warp per image row (grid-stride over the B * H rows); a lane's quads are 32 apart, so
// every store instruction of the warp writes 512 contiguous bytes.  Launched FIRST in the render
// call: it depends on nothing the other passes produce, releases its dependents at once and
// occupies a few CTAs per SM, so setup ... cover run beside it; only 4c waits for it.
__global__ void __launch_bounds__(256, 8)
raster_clear_kernel(Renderer R, RasterOut O, int n_batch) {
  pdl_launch_dependents();
  const int W = R.W, H = R.H, lane = threadIdx.x & 31;
  const int n_warps = (gridDim.x * blockDim.x) >> 5;
  const int n_rows = n_batch * H;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < n_rows; row += n_warps) {
    const int b = row / H, y = row - b * H;
    for (int x = lane * 4; x < W; x += 128)
      store_quad(O, b, H, W, y, x, z, z, z, z, make_int4(-1, -1, -1, -1));
  }
}

// 4a'.  Incremental background: the output buffers still hold this renderer's previous frame, so
// only the tiles that frame covered (the previous call's item list) need the background again.
__global__ void __launch_bounds__(256)
raster_clear_tiles_kernel(Renderer R, RasterOut O) {
  pdl_launch_dependents();
  const int n = *R.prev_heavy;
  const int sub = threadIdx.x / (RT_PIX / 4), q = threadIdx.x % (RT_PIX / 4);
  const int row4 = q / (RT_TW / 4), xq = (q % (RT_TW / 4)) * 4;
  constexpr int kPer = RT_THREADS / (RT_PIX / 4);  // tiles per block and pass
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = blockIdx.x * kPer + sub; i < n; i += gridDim.x * kPer) {
    const int work = __ldg(R.prev_list + i);
    const int y = ((work >> 8) & 0xff) * RT_TH + row4, x = (work & 0xff) * RT_TW + xq;
    if (y < R.H && x < R.W)
      store_quad(O, work >> 16, R.H, R.W, y, x, z, z, z, z, make_int4(-1, -1, -1, -1));
  }
}

// 4b.  Persistent; one non-empty (tile, pair) item per ticket.
__global__ void __launch_bounds__(RT_CV_THREADS, 8)
raster_cover_kernel(Renderer R, int n_tiles) {
  __shared__ unsigned long long zbuf[RT_PIX];
  __shared__ int s_unit[2];
  __shared__ int s_big[RT_BIG_CAP];
  __shared__ int s_nbig;
  pdl_wait();
  const int n_heavy = R.work_count[RT_WC_BINNED];
  if (threadIdx.x == 0) s_unit[0] = atomicAdd(R.work_count + RT_WC_TICKET_H, 1);
  __syncthreads();
  for (int par = 0;; par ^= 1) {
    const int unit = s_unit[par];
    if (unit >= n_heavy) break;
    if (threadIdx.x == 0) {  // next ticket, early
      s_unit[par ^ 1] = atomicAdd(R.work_count + RT_WC_TICKET_H, 1);
      s_nbig = 0;
    }
    const int work = R.bin_list[unit];
    const int b = work >> 16;
    const int tx0 = (work & 0xff) * RT_TW, ty0 = ((work >> 8) & 0xff) * RT_TH;
    const int tile = ((work >> 8) & 0xff) * R.tiles_x + (work & 0xff);
    const int32_t *start = R.tile_start + (size_t)b * (n_tiles + 1);
    const int first = start[tile];
    const int n = min(start[tile + 1], R.list_cap) - first;
    const TriRec *recs = R.recs + (size_t)b * 2 * R.max_tris;
#pragma unroll
    for (int k = 0; k < RT_PIX / RT_CV_THREADS; ++k) zbuf[threadIdx.x + k * RT_CV_THREADS] = ~0ull;
    __syncthreads();
    // ---- one thread per triangle of the tile's list
    const int32_t *list = R.tile_list + (size_t)b * R.list_cap + first;
    for (int i = threadIdx.x; i < n; i += RT_CV_THREADS) {
      const int ri = list[i];
      const int4 *rp = reinterpret_cast<const int4 *>(recs + ri);
      const int4 q0 = __ldg(rp), q1 = __ldg(rp + 1), q2 = __ldg(rp + 2), q3 = __ldg(rp + 3);
      const int xl = max(q1.w & 0xffff, tx0), xh = min(q2.x & 0xffff, tx0 + RT_TW - 1);
      const int yl = max(q1.w >> 16, ty0), yh = min(q2.x >> 16, ty0 + RT_TH - 1);
      if ((xh - xl + 1) * (yh - yl + 1) > RT_BIG_AREA) {
        const int slot = atomicAdd(&s_nbig, 1);
        if (slot < RT_BIG_CAP) {
          s_big[slot] = ri;
          continue;
        }
      }
      const int ax = q0.x, ay = q0.y, bx = q0.z, by = q0.w, cx = q1.x, cy = q1.y;
      const double zwa = __hiloint2double(q2.w, q2.z), zwb = __hiloint2double(q3.y, q3.x),
                   zwc = __hiloint2double(q3.w, q3.z);
      const int bias0 = tie_bias(cx - bx, cy - by), bias1 = tie_bias(ax - cx, ay - cy),
                bias2 = tie_bias(bx - ax, by - ay);
      const long long s0 = -(long long)(cy - by) * 256, s1 = -(long long)(ay - cy) * 256,
                      s2 = -(long long)(by - ay) * 256;
      // window depth is affine in the pixel: d = zwa + w1 * dzb + w2 * dzc (one division per
      // triangle; the same expression in the cooperative path below)
      const double inv_area = 1.0 / (double)edge_fn(ax, ay, bx, by, cx, cy);
      const double dzb = (zwb - zwa) * inv_area, dzc = (zwc - zwa) * inv_area;
      for (int y = yl; y <= yh; ++y) {
        const int px = xl << RT_SUB, py = y << RT_SUB;
        long long w0 = edge_fn(bx, by, cx, cy, px, py);
        long long w1 = edge_fn(cx, cy, ax, ay, px, py);
        long long w2 = edge_fn(ax, ay, bx, by, px, py);
        for (int x = xl; x <= xh; ++x) {
          if (((w0 + bias0) | (w1 + bias1) | (w2 + bias2)) >= 0) {
            // window depth, interpolated linearly in screen space, stored as float32
            const float d = (float)(zwa + (double)w1 * dzb + (double)w2 * dzc);
            if (d >= 0.0f && d <= 1.0f) {  // depth clipping of the fragment (far plane)
              const unsigned long long key =
                  ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)ri;  // GL_LESS
              atomicMin(&zbuf[(y - ty0) * RT_TW + (x - tx0)], key);
            }
          }
          w0 += s0;
          w1 += s1;
          w2 += s2;
        }
      }
    }
    __syncthreads();
    // ---- the queued large triangles: every thread tests its own pixels (same integer edge
    // functions, same depth expression), so no atomics are needed
    unsigned long long best[RT_PIX / RT_CV_THREADS];
#pragma unroll
    for (int k = 0; k < RT_PIX / RT_CV_THREADS; ++k) best[k] = zbuf[threadIdx.x + k * RT_CV_THREADS];
    {
      const int n_big = min(s_nbig, RT_BIG_CAP);
#pragma unroll 1
      for (int j = 0; j < n_big; ++j) {
        const int ri = s_big[j];
        const int4 *rp = reinterpret_cast<const int4 *>(recs + ri);
        const int4 q0 = __ldg(rp), q1 = __ldg(rp + 1), q2 = __ldg(rp + 2), q3 = __ldg(rp + 3);
        const int ax = q0.x, ay = q0.y, bx = q0.z, by = q0.w, cx = q1.x, cy = q1.y;
        const int bias0 = tie_bias(cx - bx, cy - by), bias1 = tie_bias(ax - cx, ay - cy),
                  bias2 = tie_bias(bx - ax, by - ay);
        const double zwa = __hiloint2double(q2.w, q2.z), zwb = __hiloint2double(q3.y, q3.x),
                     zwc = __hiloint2double(q3.w, q3.z);
        const double inv_area = 1.0 / (double)edge_fn(ax, ay, bx, by, cx, cy);
        const double dzb = (zwb - zwa) * inv_area, dzc = (zwc - zwa) * inv_area;
#pragma unroll
        for (int k = 0; k < RT_PIX / RT_CV_THREADS; ++k) {
          const int p = threadIdx.x + k * RT_CV_THREADS;
          const int px_x = tx0 + (p & (RT_TW - 1)), px_y = ty0 + (p >> RT_TW_SHIFT);
          if (px_x < (q1.w & 0xffff) || px_x > (q2.x & 0xffff) || px_y < (q1.w >> 16) ||
              px_y > (q2.x >> 16))
            continue;
          const int px = px_x << RT_SUB, py = px_y << RT_SUB;
          const long long w0 = edge_fn(bx, by, cx, cy, px, py);
          const long long w1 = edge_fn(cx, cy, ax, ay, px, py);
          const long long w2 = edge_fn(ax, ay, bx, by, px, py);
          if (((w0 + bias0) | (w1 + bias1) | (w2 + bias2)) >= 0) {
            const float d = (float)(zwa + (double)w1 * dzb + (double)w2 * dzc);
            if (d >= 0.0f && d <= 1.0f) {
              const unsigned long long key =
                  ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)ri;
              best[k] = min(best[k], key);
            }
          }
        }
      }
    }
    // merge into the pair's dense z-winner blocks, where setup's micro-triangles already are
    // (same keys, same order: the minimum is the winner whoever wrote it)
    unsigned long long *zw = R.zwin + ((size_t)b * n_tiles + tile) * RT_PIX;
#pragma unroll
    for (int k = 0; k < RT_PIX / RT_CV_THREADS; ++k)
      if (best[k] != ~0ull) atomicMin(zw + threadIdx.x + k * RT_CV_THREADS, best[k]);
    __syncthreads();  // z-buffer and queue reused by the next item; next ticket visible
  }
}

// 4c.  Grid-stride over the non-empty items, one thread per pixel, no shared memory.  Reads the
// pixel's winning key from the pair's dense z-winner blocks and puts the "empty" key back, so
// the blocks are all-empty again when the render ends (no clearing pass, no per-frame memset of
// 2 KB per tile).
template <bool LIT>
__global__ void __launch_bounds__(RT_THREADS, LIT ? 3 : 4)
raster_resolve_kernel(Renderer R, const int32_t *__restrict__ cls_idx,
                      const float *__restrict__ poses, Cam cam, RasterOut O) {
  const int W = R.W, H = R.H;
  const int n_heavy = *R.heavy;
  const int n_tiles = R.tiles_x * R.tiles_y;
  const size_t plane = (size_t)H * W;
  // A pixel's loads form a chain (item -> winning key -> vertex ids -> projected vertices, 1/z,
  // uv -> texel) and a warp has nothing else to do meanwhile: the item and the key of the NEXT
  // trip are fetched while this one shades.
  auto fetch = [&](int unit, int &work, unsigned long long *&zp, unsigned long long &key) {
    work = -1;
    zp = nullptr;
    key = ~0ull;
    if (unit >= n_heavy) return;
    work = __ldg(R.work_list + unit);
    const int x = (work & 0xff) * RT_TW + (threadIdx.x & (RT_TW - 1));
    const int y = ((work >> 8) & 0xff) * RT_TH + (threadIdx.x >> RT_TW_SHIFT);
    if (x >= W || y >= H) {
      work = -1;
      return;
    }
    const int tile = ((work >> 8) & 0xff) * R.tiles_x + (work & 0xff);
    zp = R.zwin + ((size_t)(work >> 16) * n_tiles + tile) * RT_PIX + threadIdx.x;
    key = *zp;
  };
  int work_n;
  unsigned long long *zp_n, key_n;
  fetch(blockIdx.x, work_n, zp_n, key_n);
  for (int unit = blockIdx.x; unit < n_heavy; unit += gridDim.x) {
    const int work = work_n;
    unsigned long long *zp = zp_n;
    const unsigned long long key = key_n;
    fetch(unit + gridDim.x, work_n, zp_n, key_n);
    if (work < 0) continue;  // pixel outside the frame (partial tile)
    const int b = work >> 16;
    const int x = (work & 0xff) * RT_TW + (threadIdx.x & (RT_TW - 1));
    const int y = ((work >> 8) & 0xff) * RT_TH + (threadIdx.x >> RT_TW_SHIFT);
    float dz = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
    int tid = -1;
    if (key != ~0ull) {
      *zp = ~0ull;
      // readback linearisation, render_py_multi.py:139-145, float32 like numpy
      const float d = __uint_as_float((unsigned)(key >> 32));
      dz = (d == 1.0f) ? 0.f : O.lin_c0 / (O.lin_c1 - O.lin_c2 * (2.0f * d - 1.0f));
      tid = (int)((unsigned)key >> 1);
      if (O.image_chw || O.bgr_hwc) {
        const int mesh = __ldg(cls_idx + b);
        WinTri w;
        load_winner(R, b, mesh, (unsigned)key, w);
        const Shaded s = shade_pixel<LIT>(R, O, cam, mesh, w, poses + b * 12, b, x, y);
        cr = s.c[0];
        cg = s.c[1];
        cb = s.c[2];
      }
    }
    const size_t pix = (size_t)y * W + x;
    O.depth[b * plane + pix] = dz;
    if (O.mask) O.mask[b * plane + pix] = dz > O.mask_thresh ? 1.f : 0.f;
    if (O.image_chw) {
      float *img = O.image_chw + (size_t)b * 3 * plane + pix;
      img[0] = cr - O.mean[0];
      img[plane] = cg - O.mean[1];
      img[2 * plane] = cb - O.mean[2];
    }
    if (O.bgr_hwc) {
      float *o = O.bgr_hwc + ((size_t)b * plane + pix) * 3;
      o[0] = cb;
      o[1] = cg;
      o[2] = cr;
    }
    if (O.tri_id) O.tri_id[b * plane + pix] = tid;
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
  void *ptrs[] = {r->verts, r->uv, r->normals, r->tris, r->mesh_tri_off, r->mesh_vert_off, r->vtx, r->viz,
                  r->mesh_has, r->tex,
                  r->tex_off, r->tex_hw, r->recs, r->tbox, r->clip_list, r->clip_box, r->zwin, r->work_count, r->tile_start, r->tile_fill,
                  r->tile_list, r->work_list_b[0], r->work_list_b[1], r->heavy_count, r->bin_list, r->bin_tri, r->bin_n};
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
  RC_REQUIRE(width % 4 == 0, RC_ERR_UNALIGNED);
  // work items pack (pair, tile row, tile column) into 15 + 8 + 8 bits
  RC_REQUIRE(width <= 255 * rc::RT_TW && height <= 255 * rc::RT_TH && max_batch <= 32767,
             RC_ERR_CAPACITY);
  size_t V = 0, T = 0, TX = 0;
  int max_tris = 0, max_verts = 0;
  for (int m = 0; m < n_meshes; ++m) {
    const rc_mesh_desc &d = meshes[m];
    RC_REQUIRE(d.vertices && d.triangles && d.n_vertices > 0 && d.n_triangles > 0,
               RC_ERR_INVALID_ARG);
    RC_REQUIRE(d.texture && d.tex_h > 0 && d.tex_w > 0, RC_ERR_INVALID_ARG);
    V += d.n_vertices;
    T += d.n_triangles;
    TX += (size_t)d.tex_h * d.tex_w * 3;
    if (d.n_triangles > max_tris) max_tris = d.n_triangles;
    if (d.n_vertices > max_verts) max_verts = d.n_vertices;
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
  r->max_verts = max_verts;
  r->tiles_x = (width + rc::RT_TW - 1) / rc::RT_TW;
  r->tiles_y = (height + rc::RT_TH - 1) / rc::RT_TH;
  const int n_tiles = r->tiles_x * r->tiles_y;
  r->list_cap = 8 * max_tris + 16 * n_tiles;
  // host staging of the concatenated mesh arrays
  std::vector<float> hv(V * 3), huv(V * 2, 0.f), hn(V * 3, 0.f);
  std::vector<int32_t> ht(T * 3), hoff(n_meshes + 1), hvoff(n_meshes + 1), hhas(n_meshes),
      hhw(n_meshes * 2);
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
    hvoff[m] = (int32_t)v0;
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
  hvoff[n_meshes] = (int32_t)v0;
  RC_TRY(rc::dev_alloc(&r->verts, V * 3));
  RC_TRY(rc::dev_alloc(&r->uv, V * 2));
  RC_TRY(rc::dev_alloc(&r->normals, V * 3));
  RC_TRY(rc::dev_alloc(&r->tris, T * 3));
  RC_TRY(rc::dev_alloc(&r->mesh_tri_off, (size_t)n_meshes + 1));
  RC_TRY(rc::dev_alloc(&r->mesh_vert_off, (size_t)n_meshes + 1));
  RC_TRY(rc::dev_alloc(&r->vtx, (size_t)max_batch * max_verts));
  RC_TRY(rc::dev_alloc(&r->viz, (size_t)max_batch * max_verts));
  RC_TRY(rc::dev_alloc(&r->mesh_has, (size_t)n_meshes));
  RC_TRY(rc::dev_alloc(&r->tex, TX));
  RC_TRY(rc::dev_alloc(&r->tex_off, (size_t)n_meshes));
  RC_TRY(rc::dev_alloc(&r->tex_hw, (size_t)n_meshes * 2));
  RC_TRY(rc::dev_alloc(&r->recs, (size_t)max_batch * 2 * max_tris));
  RC_TRY(rc::dev_alloc(&r->tbox, (size_t)max_batch * max_tris));
  RC_TRY(rc::dev_alloc(&r->clip_list, (size_t)max_batch * max_tris));
  RC_TRY(rc::dev_alloc(&r->clip_box, (size_t)max_batch * max_tris));
  // [header | tile counts of the call's B pairs | micro-triangle marks of the same pairs]
  RC_TRY(rc::dev_alloc(&r->work_count, (size_t)2 * max_batch * n_tiles + rc::RT_WC_INTS));
  r->tile_count = r->work_count + rc::RT_WC_INTS;
  RC_TRY(rc::dev_alloc(&r->bin_list, (size_t)max_batch * n_tiles));
  RC_TRY(rc::dev_alloc(&r->bin_tri, (size_t)max_batch * max_tris));
  RC_TRY(rc::dev_alloc(&r->bin_n, (size_t)max_batch));
  RC_TRY(rc::dev_alloc(&r->work_list_b[0], (size_t)max_batch * n_tiles));
  RC_TRY(rc::dev_alloc(&r->work_list_b[1], (size_t)max_batch * n_tiles));
  RC_TRY(rc::dev_alloc(&r->heavy_count, 2));
  RC_TRY(rc::dev_alloc(&r->zwin, (size_t)max_batch * n_tiles * rc::RT_PIX));
  // all-empty; every render leaves it so (resolve puts the empty key back where it read a winner)
  if (cudaMemset(r->zwin, 0xff, (size_t)max_batch * n_tiles * rc::RT_PIX * sizeof(unsigned long long)) !=
          cudaSuccess ||
      cudaMemset(r->heavy_count, 0, 2 * sizeof(int32_t)) != cudaSuccess) {
    rc::renderer_free(r);
    return rc::note_status(RC_ERR_CUDA);
  }
  RC_TRY(rc::dev_alloc(&r->tile_start, (size_t)max_batch * (n_tiles + 1)));
  RC_TRY(rc::dev_alloc(&r->tile_fill, (size_t)max_batch * n_tiles));
  RC_TRY(rc::dev_alloc(&r->tile_list, (size_t)max_batch * r->list_cap));
#define RC_UP(dst, src) \
  if (cudaMemcpy(dst, src.data(), src.size() * sizeof(src[0]), cudaMemcpyHostToDevice) != cudaSuccess) { \
    rc::renderer_free(r); \
    return rc::note_status(RC_ERR_CUDA); \
  }
  RC_UP(r->verts, hv) RC_UP(r->uv, huv) RC_UP(r->normals, hn) RC_UP(r->tris, ht)
  RC_UP(r->mesh_tri_off, hoff) RC_UP(r->mesh_vert_off, hvoff) RC_UP(r->mesh_has, hhas)
  RC_UP(r->tex, htex)
  RC_UP(r->tex_off, htoff) RC_UP(r->tex_hw, hhw)
#undef RC_UP
  // the memsets / copies above ran on the legacy default stream; the renders will come on the
  // caller's (possibly non-blocking) streams
  if (cudaDeviceSynchronize() != cudaSuccess) {
    rc::renderer_free(r);
    return rc::note_status(RC_ERR_CUDA);
  }
  *handle = r;
  return rc::note_status(RC_OK);
}

int rc_renderer_destroy(void *handle) {
  rc::renderer_free(static_cast<rc::Renderer *>(handle));
  return rc::note_status(RC_OK);
}

int rc_renderer_note_replay(void *handle, int n_renders) {
  RC_REQUIRE(handle != nullptr && n_renders >= 0, RC_ERR_INVALID_ARG);
  rc::Renderer &Rh = *static_cast<rc::Renderer *>(handle);
  Rh.cur ^= (n_renders & 1);
  return rc::note_status(RC_OK);
}

int rc_render_dev(void *handle, const rc_render_args *a) {
  RC_REQUIRE(handle && a, RC_ERR_INVALID_ARG);
  rc::Renderer &Rh = *static_cast<rc::Renderer *>(handle);  // the handle (host state lives here)
  rc::Renderer R = Rh;                                       // this call's view, passed to kernels
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
      RC_CUDA_TRY(rc::record_mark(static_cast<cudaEvent_t>(a->marks[i]), s));
    return RC_OK;
  };
  int st;
  const int sm_count = rc::device_sm_count(R.device);
  static int cover_per_sm = 0, resolve_per_sm[2] = {0, 0};  // per kernel, not per device
  if (cover_per_sm == 0) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cover_per_sm, rc::raster_cover_kernel,
                                                      rc::RT_CV_THREADS, 0) != cudaSuccess ||
        cover_per_sm < 1)
      cover_per_sm = 4;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resolve_per_sm[0],
                                                      rc::raster_resolve_kernel<false>,
                                                      rc::RT_THREADS, 0) != cudaSuccess ||
        resolve_per_sm[0] < 1)
      resolve_per_sm[0] = 2;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resolve_per_sm[1],
                                                      rc::raster_resolve_kernel<true>,
                                                      rc::RT_THREADS, 0) != cudaSuccess ||
        resolve_per_sm[1] < 1)
      resolve_per_sm[1] = 2;
  }
  const long long n_total = (long long)n_tiles * B;
  // With event marks every launch is a normal one (clean per-kernel times); without, setup ...
  // cover are chained by programmatic dependent launches behind the background pass: each waits
  // (griddepcontrol.wait) for its predecessor only, none for the clear, which they never touch.
  const bool chain = a->marks == nullptr;
  if ((st = mark(0)) != RC_OK) return st;
  // ping-pong item lists: this call fills one half, the other half still names the tiles the
  // previous frame dirtied
  const int cur = Rh.cur;
  R.work_list = Rh.work_list_b[cur];
  R.heavy = Rh.heavy_count + cur;
  R.prev_list = Rh.work_list_b[cur ^ 1];
  R.prev_heavy = Rh.heavy_count + (cur ^ 1);
  const void *outs[5] = {a->depth, a->image_chw, a->bgr_hwc, a->mask, a->tri_id};
  bool same = Rh.have_prev && Rh.last_B == B;
  for (int i = 0; i < 5 && same; ++i) same = Rh.last_out[i] == outs[i];
  for (int i = 0; i < 3 && same; ++i) same = Rh.last_mean[i] == a->pixel_means[i];
  const bool incremental = a->incremental != 0 && same;
  Rh.have_prev = 0;  // until this call's launches are all issued
  // (no memsets: the vertex pass zeroes this call's counters)
  R.tile_touch = R.tile_count + (size_t)B * n_tiles;
  // 4a background first: a few CTAs per SM saturate the HBM write path and leave room beside them
  static int clear_ctas = 0;
  if (clear_ctas == 0) {
    const char *e = getenv("RC_CLEAR_CTAS_PER_SM");  // tuning knob; default below
    clear_ctas = e ? atoi(e) : 2;
    if (clear_ctas < 1 || clear_ctas > 8) clear_ctas = 2;
  }
  if (incremental)
    rc::raster_clear_tiles_kernel<<<sm_count * clear_ctas, 256, 0, s>>>(R, O);
  else
    rc::raster_clear_kernel<<<sm_count * clear_ctas, 256, 0, s>>>(R, O, B);
  dim3 gv((R.max_verts + 255) / 256, B);
  RC_CUDA_TRY(rc::launch_maybe_pdl(rc::raster_vertex_kernel, gv, dim3(256), 0, s, chain, R,
                                   a->cls_idx, a->poses, cam, n_tiles));
  const int per_block = 256 * rc::RT_SETUP_PER;
  dim3 gs((R.max_tris + per_block - 1) / per_block, B);
  RC_CUDA_TRY(rc::launch_maybe_pdl(rc::raster_setup_kernel, gs, dim3(256), 0, s, chain, R,
                                   a->cls_idx, n_tiles));
  if ((st = mark(1)) != RC_OK) return st;
  RC_CUDA_TRY(rc::launch_maybe_pdl(rc::raster_scan_kernel, dim3(B), dim3(256), 0, s, chain, R,
                                   a->cls_idx, a->poses, cam, n_tiles, a->status));
  if ((st = mark(2)) != RC_OK) return st;
  dim3 gf((unsigned)std::min(8, (R.max_tris + 255) / 256), B);
  RC_CUDA_TRY(rc::launch_maybe_pdl(rc::raster_fill_kernel, gf, dim3(256), 0, s, chain, R,
                                   a->cls_idx, n_tiles));
  if ((st = mark(3)) != RC_OK) return st;
  {
    long long g = (long long)sm_count * cover_per_sm;
    if (g > n_total) g = n_total;
    RC_CUDA_TRY(rc::launch_maybe_pdl(rc::raster_cover_kernel, dim3((unsigned)g),
                                     dim3(rc::RT_CV_THREADS), 0, s, chain, R, n_tiles));
  }
  // 4c resolve: a normal launch, waits for everything before it, the background included
  {
    long long g = (long long)sm_count * resolve_per_sm[a->mode == RC_RENDER_TEXTURED ? 0 : 1] * 2;
    if (g > n_total) g = n_total;
    if (a->mode == RC_RENDER_TEXTURED)
      rc::raster_resolve_kernel<false><<<(int)g, rc::RT_THREADS, 0, s>>>(R, a->cls_idx, a->poses, cam, O);
    else
      rc::raster_resolve_kernel<true><<<(int)g, rc::RT_THREADS, 0, s>>>(R, a->cls_idx, a->poses, cam, O);
  }
  if ((st = mark(4)) != RC_OK) return st;
  st = rc::check_launch();
  // host state for the next call's incremental test (the handle is not thread-safe)
  Rh.have_prev = st == RC_OK;
  Rh.n_frames = st == RC_OK ? (Rh.n_frames < 2 ? Rh.n_frames + 1 : 2) : 0;
  Rh.last_B = B;
  for (int i = 0; i < 5; ++i) Rh.last_out[i] = outs[i];
  for (int i = 0; i < 3; ++i) Rh.last_mean[i] = a->pixel_means[i];
  Rh.cur = cur ^ 1;
  return rc::note_status(st);
}

}  // extern "C"

namespace rc {
// The last frame's items for the compare step's sparse passes (frame.cu).  Valid until the
// next rc_render_dev call on this handle.
int renderer_frame_items(void *handle, FrameItems *out) {
  RC_REQUIRE(handle && out, RC_ERR_INVALID_ARG);
  const Renderer &R = *static_cast<const Renderer *>(handle);
  RC_REQUIRE(R.have_prev && R.n_frames >= 1, RC_ERR_INVALID_ARG);  // no completed render yet
  const int last = R.cur ^ 1;  // the half the last call filled
  out->cur_list = R.work_list_b[last];
  out->cur_count = R.heavy_count + last;
  const bool two = R.n_frames >= 2;
  out->prev_list = two ? R.work_list_b[last ^ 1] : nullptr;
  out->prev_count = two ? R.heavy_count + (last ^ 1) : nullptr;
  out->tile_count = R.tile_count;
  out->tiles_x = R.tiles_x;
  out->tiles_y = R.tiles_y;
  out->H = R.H;
  out->W = R.W;
  out->max_batch = R.max_batch;
  return RC_OK;
}
}  // namespace rc

extern "C" {

int rc_render_stats(void *handle, int batch_size, void *stream, int64_t *out) {
  RC_REQUIRE(handle && out && batch_size > 0, RC_ERR_INVALID_ARG);
  rc::Renderer &R = *static_cast<rc::Renderer *>(handle);
  RC_REQUIRE(batch_size <= R.max_batch, RC_ERR_CAPACITY);
  const int n_tiles = R.tiles_x * R.tiles_y;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  RC_CUDA_TRY(cudaStreamSynchronize(s));
  std::vector<int32_t> st((size_t)batch_size * (n_tiles + 1));
  RC_CUDA_TRY(cudaMemcpy(st.data(), R.tile_start, st.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
  std::vector<uint32_t> bb((size_t)batch_size * R.max_tris);
  RC_CUDA_TRY(cudaMemcpy(bb.data(), R.tbox, bb.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  int32_t n_clip = 0;
  RC_CUDA_TRY(cudaMemcpy(&n_clip, R.work_count + rc::RT_WC_CLIP, sizeof(int32_t), cudaMemcpyDeviceToHost));
  std::vector<uint2> cb((size_t)n_clip);
  if (n_clip > 0)
    RC_CUDA_TRY(cudaMemcpy(cb.data(), R.clip_box, cb.size() * sizeof(uint2), cudaMemcpyDeviceToHost));
  int64_t recs = 0, entries = 0;
  for (int b = 0; b < batch_size; ++b) entries += st[(size_t)b * (n_tiles + 1) + n_tiles];
  // live records (slots of absent triangles count only when the mesh has max_tris triangles:
  // stale boxes beyond a smaller mesh's triangles are never read by the passes)
  // (micro-triangles that went straight to the z-winner blocks count as records: out[2] = how many)
  int64_t direct = 0;
  for (size_t i = 0; i < bb.size(); ++i) {
    recs += bb[i] != rc::RT_TBOX_NONE;
    direct += bb[i] == rc::RT_TBOX_DIRECT;
  }
  for (size_t i = 0; i < cb.size(); ++i)
    recs += (cb[i].x != rc::RT_TBOX_NONE) + (cb[i].y != rc::RT_TBOX_NONE);
  out[0] = recs;
  out[1] = entries;
  out[2] = direct;
  return rc::note_status(RC_OK);
}

}  // extern "C"
