// eval.cu -- evaluation reductions on the step after the path:
//
//   /root/reference/lib/utils/pose_error.py:55-147   arp_2d, add, adi, re, te
//   /root/reference/lib/utils/visibility.py:7-33     estimate_visib_mask(_gt/_est)
//   /root/reference/deepim/core/tester.py:719-736    calc_EPE_one_pair
//
// All per-point / per-pixel reductions over buffers the path already holds on the device.
// Sums are float64 and deterministic: every block writes its partial sums to a workspace slot
// and a one-block-per-pair finishing kernel adds the slots in index order (no float atomics).
#include "common.cuh"

namespace rc {

__device__ __forceinline__ double block_sum_256(double v, double *red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
  }
  return t;  // valid in thread 0
}

__device__ __forceinline__ void xform(const float *__restrict__ P, double x, double y, double z,
                                      double (&o)[3]) {
#pragma unroll
  for (int r = 0; r < 3; ++r)
    o[r] = fma((double)P[r * 4 + 2], z, fma((double)P[r * 4 + 1], y, (double)P[r * 4] * x)) +
           (double)P[r * 4 + 3];
}

constexpr int PE_THREADS = 256;
constexpr int PE_TILE = 1024;  // estimated points staged per pass (24 KB of float64)

// grid = (chunks, B): a block owns up to 256 model points of one pair: their ADD and ARP-2D
// terms, and for ADI the distance from each of its GT-transformed points to the nearest of ALL
// estimate-transformed points (exact brute force, tiled through shared memory; the reference
// uses a k-d tree, which returns the same minimum).
//   partial [B, chunks, 3] : sums of add / adi / arp terms of the chunk
__global__ void __launch_bounds__(PE_THREADS)
pose_error_kernel(const float *__restrict__ pose_est, const float *__restrict__ pose_gt,
                  const float *__restrict__ points, long long points_bstride,
                  const int32_t *__restrict__ n_valid, const float *__restrict__ K, int N,
                  int want_adi, double *__restrict__ partial) {
  __shared__ double tile[3][PE_TILE];
  __shared__ double red[PE_THREADS / 32];
  const int b = blockIdx.y;
  const int n = n_valid ? min(max(n_valid[b], 0), N) : N;
  const float *pts = points + (long long)b * points_bstride;
  const float *Pe = pose_est + b * 12, *Pg = pose_gt + b * 12;
  const int i = blockIdx.x * PE_THREADS + threadIdx.x;
  const bool act = i < n;
  double g[3] = {0, 0, 0}, e[3] = {0, 0, 0};
  double add_t = 0.0, arp_t = 0.0;
  if (act) {
    const double x = pts[i], y = pts[N + i], z = pts[2 * N + i];
    xform(Pe, x, y, z, e);
    xform(Pg, x, y, z, g);
    add_t = sqrt((e[0] - g[0]) * (e[0] - g[0]) + (e[1] - g[1]) * (e[1] - g[1]) +
                 (e[2] - g[2]) * (e[2] - g[2]));
    // transform_pts_Rt_2d: K . X, divide by the third row
    double pe[3], pg[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      pe[r] = fma((double)K[r * 3 + 2], e[2], fma((double)K[r * 3 + 1], e[1], (double)K[r * 3] * e[0]));
      pg[r] = fma((double)K[r * 3 + 2], g[2], fma((double)K[r * 3 + 1], g[1], (double)K[r * 3] * g[0]));
    }
    const double du = pe[0] / pe[2] - pg[0] / pg[2], dv = pe[1] / pe[2] - pg[1] / pg[2];
    arp_t = sqrt(du * du + dv * dv);
  }
  double best = 1.0e300;
  if (want_adi) {
    for (int base = 0; base < n; base += PE_TILE) {
      __syncthreads();
      for (int k = threadIdx.x; k < PE_TILE; k += PE_THREADS) {
        const int j = base + k;
        double q[3] = {1.0e150, 1.0e150, 1.0e150};
        if (j < n) xform(Pe, (double)pts[j], (double)pts[N + j], (double)pts[2 * N + j], q);
        tile[0][k] = q[0];
        tile[1][k] = q[1];
        tile[2][k] = q[2];
      }
      __syncthreads();
      if (act) {
        const int m = min(PE_TILE, n - base);
#pragma unroll 4
        for (int k = 0; k < m; ++k) {
          const double dx = tile[0][k] - g[0], dy = tile[1][k] - g[1], dz = tile[2][k] - g[2];
          best = fmin(best, fma(dz, dz, fma(dy, dy, dx * dx)));
        }
      }
    }
  }
  const double adi_t = (act && want_adi) ? sqrt(best) : 0.0;
  double *out = partial + ((size_t)b * gridDim.x + blockIdx.x) * 3;
  const double s0 = block_sum_256(add_t, red);
  const double s1 = block_sum_256(adi_t, red);
  const double s2 = block_sum_256(arp_t, red);
  if (threadIdx.x == 0) {
    out[0] = s0;
    out[1] = s1;
    out[2] = s2;
  }
}

// One thread per pair: means of the partial sums, rotation error (the angle of R_est^T R_gt,
// which is what ||logm(.)||_F / sqrt(2) evaluates to, pose_error.py:126-131) and translation
// error (:134-145).  out [B,5] float64: add, adi, arp_2d, re (deg), te.
__global__ void pose_error_finish_kernel(const float *__restrict__ pose_est,
                                         const float *__restrict__ pose_gt,
                                         const int32_t *__restrict__ n_valid,
                                         const double *__restrict__ partial, int chunks, int N,
                                         int B, double *__restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int n = n_valid ? min(max(n_valid[b], 0), N) : N;
  double s[3] = {0, 0, 0};
  for (int c = 0; c < chunks; ++c)
#pragma unroll
    for (int k = 0; k < 3; ++k) s[k] += partial[((size_t)b * chunks + c) * 3 + k];
  const double inv = n > 0 ? 1.0 / (double)n : 0.0;
  const float *Pe = pose_est + b * 12, *Pg = pose_gt + b * 12;
  double M[3][3];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c)
      M[r][c] = (double)Pe[0 * 4 + r] * Pg[0 * 4 + c] + (double)Pe[1 * 4 + r] * Pg[1 * 4 + c] +
                (double)Pe[2 * 4 + r] * Pg[2 * 4 + c];
  const double cs = 0.5 * (M[0][0] + M[1][1] + M[2][2] - 1.0);
  const double vx = 0.5 * (M[2][1] - M[1][2]), vy = 0.5 * (M[0][2] - M[2][0]),
               vz = 0.5 * (M[1][0] - M[0][1]);
  const double sn = sqrt(vx * vx + vy * vy + vz * vz);
  const double re = atan2(sn, cs) / 3.14159265358979323846 * 180.0;
  const double tx = (double)Pg[3] - (double)Pe[3], ty = (double)Pg[7] - (double)Pe[7],
               tz = (double)Pg[11] - (double)Pe[11];
  out[b * 5 + 0] = s[0] * inv;
  out[b * 5 + 1] = s[1] * inv;
  out[b * 5 + 2] = s[2] * inv;
  out[b * 5 + 3] = re;
  out[b * 5 + 4] = sqrt(tx * tx + ty * ty + tz * tz);
}

}  // namespace rc

namespace rc {

// estimate_visib_mask (visibility.py:7-22) and estimate_visib_mask_est (:30-33), 4 px/thread:
//   visib = d_test > 0 && d_model > 0 && (d_model - d_test) <= delta          (float32)
//   with visib_gt: visib |= visib_gt && d_model > 0
__global__ void __launch_bounds__(256)
visib_mask_kernel(const float *__restrict__ d_test, const float *__restrict__ d_model,
                  const float *__restrict__ visib_gt, float delta, float *__restrict__ out,
                  long long quads) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += stride) {
    const float4 t = ld_stream_f4(reinterpret_cast<const float4 *>(d_test) + q);
    const float4 m = ld_stream_f4(reinterpret_cast<const float4 *>(d_model) + q);
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (visib_gt) g = ld_stream_f4(reinterpret_cast<const float4 *>(visib_gt) + q);
    auto one = [&](float tt, float mm, float gg) -> float {
      bool v = (tt > 0.f) && (mm > 0.f) && (__fsub_rn(mm, tt) <= delta);
      v = v || ((gg != 0.f) && (mm > 0.f));
      return v ? 1.f : 0.f;
    };
    st_f4(reinterpret_cast<float4 *>(out) + q,
          make_float4(one(t.x, m.x, g.x), one(t.y, m.y, g.y), one(t.z, m.z, g.z), one(t.w, m.w, g.w)));
  }
}

// calc_EPE_one_pair (tester.py:719-736).  grid = (chunks, B); partial [B, chunks, 5]:
// epe_all, epe_viz, num_viz, epe_vizbg, num_vizbg.  bg = visible == 0 && depth_rendered == 0
// (tester.py:712) when `bg` is NULL and depth_rendered is given.
__global__ void __launch_bounds__(256)
flow_epe_kernel(const float *__restrict__ pred, const float *__restrict__ gt,
                const float *__restrict__ visible, const float *__restrict__ bg,
                const float *__restrict__ depth_rendered, int plane, double *__restrict__ partial) {
  __shared__ double red[8];
  const int b = blockIdx.y;
  const float *p0 = pred + (size_t)b * 2 * plane, *g0 = gt + (size_t)b * 2 * plane;
  const float *vis = visible + (size_t)b * plane;
  double s[5] = {0, 0, 0, 0, 0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += gridDim.x * blockDim.x) {
    const double dx = (double)g0[i] - (double)p0[i];
    const double dy = (double)g0[plane + i] - (double)p0[plane + i];
    const double d = sqrt(dx * dx + dy * dy);
    const bool v = vis[i] == 1.f;
    bool isbg = false;
    if (bg) isbg = bg[(size_t)b * plane + i] != 0.f;
    else if (depth_rendered) isbg = (vis[i] == 0.f) && (depth_rendered[(size_t)b * plane + i] == 0.f);
    const bool vb = (vis[i] != 0.f) || isbg;  // np.logical_or(visible, bg)
    s[0] += d;
    if (v) s[1] += d;
    s[2] += (double)vis[i];
    if (vb) {
      s[3] += d;
      s[4] += 1.0;
    }
  }
  double *out = partial + ((size_t)b * gridDim.x + blockIdx.x) * 5;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double t = block_sum_256(s[k], red);
    if (threadIdx.x == 0) out[k] = t;
  }
}

__global__ void flow_epe_finish_kernel(const double *__restrict__ partial, int chunks, int plane,
                                       int B, double *__restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double s[5] = {0, 0, 0, 0, 0};
  for (int c = 0; c < chunks; ++c)
#pragma unroll
    for (int k = 0; k < 5; ++k) s[k] += partial[((size_t)b * chunks + c) * 5 + k];
  out[b * 6 + 0] = s[0];
  out[b * 6 + 1] = (double)plane;
  out[b * 6 + 2] = s[1];
  out[b * 6 + 3] = s[2];
  out[b * 6 + 4] = s[3];
  out[b * 6 + 5] = s[4];
}

}  // namespace rc

extern "C" {

int rc_pose_error_chunks(int num_points) { return (num_points + rc::PE_THREADS - 1) / rc::PE_THREADS; }

int rc_pose_error_dev(const float *pose_est, const float *pose_gt, const float *points,
                      int points_shared, const int32_t *n_valid, const float *K, int want_adi,
                      double *out, double *workspace, int batch_size, int num_points,
                      void *stream) {
  RC_REQUIRE(batch_size >= 0 && num_points > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(pose_est && pose_gt && points && K && out && workspace, RC_ERR_INVALID_ARG);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int chunks = rc_pose_error_chunks(num_points);
  dim3 grid(chunks, batch_size);
  rc::pose_error_kernel<<<grid, rc::PE_THREADS, 0, s>>>(
      pose_est, pose_gt, points, points_shared ? 0ll : 3ll * num_points, n_valid, K, num_points,
      want_adi ? 1 : 0, workspace);
  RC_CUDA_TRY(cudaGetLastError());
  rc::pose_error_finish_kernel<<<(batch_size + 63) / 64, 64, 0, s>>>(
      pose_est, pose_gt, n_valid, workspace, chunks, num_points, batch_size, out);
  return rc::note_status(rc::check_launch());
}

int rc_visib_mask_dev(const float *d_test, const float *d_model, const float *visib_gt,
                      float delta, float *out, int64_t count, void *stream) {
  RC_REQUIRE(count >= 0, RC_ERR_INVALID_ARG);
  if (count == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(d_test && d_model && out, RC_ERR_INVALID_ARG);
  RC_REQUIRE(count % 4 == 0 && rc::aligned16(d_test) && rc::aligned16(d_model) &&
                 rc::aligned16(out) && rc::aligned16(visib_gt),
             RC_ERR_UNALIGNED);
  const long long quads = count / 4;
  long long blocks = (quads + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  rc::visib_mask_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_test, d_model, visib_gt, delta, out, quads);
  return rc::note_status(rc::check_launch());
}

int rc_flow_epe_chunks(void) { return 64; }

int rc_flow_epe_dev(const float *flow_pred, const float *flow_gt, const float *visible,
                    const float *bg, const float *depth_rendered, double *out, double *workspace,
                    int batch_size, int height, int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(flow_pred && flow_gt && visible && out && workspace, RC_ERR_INVALID_ARG);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int chunks = rc_flow_epe_chunks();
  dim3 grid(chunks, batch_size);
  rc::flow_epe_kernel<<<grid, 256, 0, s>>>(flow_pred, flow_gt, visible, bg, depth_rendered,
                                           height * width, workspace);
  RC_CUDA_TRY(cudaGetLastError());
  rc::flow_epe_finish_kernel<<<(batch_size + 63) / 64, 64, 0, s>>>(workspace, chunks,
                                                                   height * width, batch_size, out);
  return rc::note_status(rc::check_launch());
}

}  // extern "C"
