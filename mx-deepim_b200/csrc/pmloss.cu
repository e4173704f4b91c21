// pmloss.cu -- Transform3D (rigid transform of model points by the predicted
// pose update) and the point-matching loss, forward and backward.
//
// Replaces /root/reference/deepim/operator_py/transform3d.py:42-118 (forward),
// :120-191 (backward), :193-225 (T_transform_backward), :227-327
// (quat2mat forward/backward) and the loss graph
// /root/reference/deepim/symbols/deepIM_flownet.py:460-499.
//
// B200 design: the reference runs three per-sample Python loops with a
// device->host sync each plus two cuBLAS batched GEMMs with a 3-wide inner
// dimension.  Here one thread-block CLUSTER of eight CTAs owns one sample
// (2048 threads for N = 3000 points, so every thread has at most two points
// and all its loads are in flight at once): thread 0 of each CTA evaluates the
// pose algebra (a few dozen scalar operations, float64 where numpy promotes to
// float64) into shared memory, all threads stream the [3,N] point arrays with
// coalesced loads, and the 1 (loss), 3 (sum G) and 9 (G X^T) sums are reduced
// with warp shuffles, one shared-memory stage per CTA and one distributed-
// shared-memory stage across the cluster (rank order) -- deterministic, no
// atomics, no global workspace.  The path moves 0.144 MB (fwd) / 0.108 MB
// (bwd) per pair, so it is latency bound; the cluster spreads a batch of 16
// pairs over 128 SMs instead of 16.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace rc {

struct PoseParams {
  float means[3];
  float stds[3];
  int rot_coord;
};

struct PoseOut {
  float R[9];      // Rm_tgt
  float T[3];      // T_tgt
  float Rd[9];     // Rm_delta (cached for backward in the reference, :81)
  float Td1[3];    // T_delta * std + mean
  float z2;
};

// quat2mat_forward, transform3d.py:227-254.  Nq in float32, the rest in float64
// (2.0 / np.float32 promotes to float64 under the reference's numpy), result
// stored float32.
__device__ void quat2mat_fwd(const float *q, float *M) {
  const float w = q[0], x = q[1], y = q[2], z = q[3];
  const float Nq = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w, w), __fmul_rn(x, x)), __fmul_rn(y, y)),
                             __fmul_rn(z, z));
  const double dn = (double)Nq - 1.0;
  if (!(-1e-2 < dn && dn < 1e-2)) {  // :230-231
    M[0] = 1.f; M[1] = 0.f; M[2] = 0.f;
    M[3] = 0.f; M[4] = 1.f; M[5] = 0.f;
    M[6] = 0.f; M[7] = 0.f; M[8] = 1.f;
    return;
  }
  const double s = 2.0 / (double)Nq;
  const double X = x * s, Y = y * s, Z = z * s;
  const double wX = w * X, wY = w * Y, wZ = w * Z;
  const double xX = x * X, xY = x * Y, xZ = x * Z;
  const double yY = y * Y, yZ = y * Z, zZ = z * Z;
  M[0] = (float)(1.0 - (yY + zZ)); M[1] = (float)(xY - wZ);         M[2] = (float)(xZ + wY);
  M[3] = (float)(xY + wZ);         M[4] = (float)(1.0 - (xX + zZ)); M[5] = (float)(yZ - wX);
  M[6] = (float)(xZ - wY);         M[7] = (float)(yZ + wX);         M[8] = (float)(1.0 - (xX + yY));
}

__device__ __forceinline__ void mat3_mul(const float *A, const float *B, float *C) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C[i * 3 + j] = __fadd_rn(__fadd_rn(__fmul_rn(A[i * 3 + 0], B[0 * 3 + j]),
                                         __fmul_rn(A[i * 3 + 1], B[1 * 3 + j])),
                               __fmul_rn(A[i * 3 + 2], B[2 * 3 + j]));
}

// Pose composition of transform3d.py:70-112 with R_transform / T_transform
// from lib/pair_matching/RT_transform.py:51-110.
__device__ void compose_pose(const float *q, const float *td, const float *pose,
                             const PoseParams &pp, PoseOut &o) {
  float Rs[9], Ts[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    Rs[i * 3 + 0] = pose[i * 4 + 0];
    Rs[i * 3 + 1] = pose[i * 4 + 1];
    Rs[i * 3 + 2] = pose[i * 4 + 2];
    Ts[i] = pose[i * 4 + 3];
  }
  quat2mat_fwd(q, o.Rd);
  if (pp.rot_coord == RC_ROT_MODEL) mat3_mul(Rs, o.Rd, o.R);  // RT_transform.py:59-60
  else mat3_mul(o.Rd, Rs, o.R);                               // :61-66
#pragma unroll
  for (int i = 0; i < 3; ++i) o.Td1[i] = __fadd_rn(__fmul_rn(td[i], pp.stds[i]), pp.means[i]);
  if (pp.rot_coord == RC_ROT_NAIVE) {  // T_transform_naive, RT_transform.py:106-110
#pragma unroll
    for (int i = 0; i < 3; ++i)
      o.T[i] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(o.Rd[i * 3 + 0], Ts[0]),
                                             __fmul_rn(o.Rd[i * 3 + 1], Ts[1])),
                                   __fmul_rn(o.Rd[i * 3 + 2], Ts[2])),
                         td[i]);
    o.z2 = o.T[2];
    return;
  }
  const float z2 = __fdiv_rn(Ts[2], expf(o.Td1[2]));  // :92
  o.z2 = z2;
  o.T[2] = z2;
  if (pp.rot_coord == RC_ROT_CAMERA_NEW) {  // :97-99
    o.T[0] = __fadd_rn(__fmul_rn(Ts[2], o.Td1[0]), Ts[0]);
    o.T[1] = __fadd_rn(__fmul_rn(Ts[2], o.Td1[1]), Ts[1]);
  } else {  // camera / model, :94-96
    o.T[0] = __fmul_rn(z2, __fadd_rn(o.Td1[0], __fdiv_rn(Ts[0], Ts[2])));
    o.T[1] = __fmul_rn(z2, __fadd_rn(o.Td1[1], __fdiv_rn(Ts[1], Ts[2])));
  }
}

__device__ __forceinline__ float loss_value(float x, int type, float sigma2, float inv_sigma2) {
  if (type == RC_PM_L1) return fabsf(x);
  if (type == RC_PM_L2) return __fmul_rn(x, x);
  // mx.sym.smooth_l1(scalar=sigma): 0.5 (sigma x)^2 if |x| < 1/sigma^2 else |x| - 0.5/sigma^2
  const float ax = fabsf(x);
  return ax < inv_sigma2 ? 0.5f * sigma2 * x * x : ax - 0.5f * inv_sigma2;
}

__device__ __forceinline__ float loss_grad(float x, int type, float sigma2, float inv_sigma2) {
  const float sgn = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
  if (type == RC_PM_L1) return sgn;
  if (type == RC_PM_L2) return 2.f * x;
  return fabsf(x) < inv_sigma2 ? sigma2 * x : sgn;
}

constexpr int PM_THREADS = 256;
constexpr int PM_CLUSTER = 8;  // CTAs per sample at small batches (portable cluster size limit)

// Sum NV per-thread values over the block in ONE pass: warp shuffles, one shared-memory stage
// (sm [PM_THREADS / 32][NV]) between two barriers, warps added in index order by thread 0.
// The result is valid in thread 0.
template <int NV>
__device__ __forceinline__ void block_sum_n(const double (&v)[NV], double *sm, double (&out)[NV]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double w[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) w[i] = warp_sum(v[i]);
  __syncthreads();  // sm may still be read by a previous call
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) sm[warp * NV + i] = w[i];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      double r = 0.0;
#pragma unroll
      for (int k = 0; k < PM_THREADS / 32; ++k) r += sm[k * NV + i];
      out[i] = r;
    }
  }
}

// Sum NV per-CTA partials across the cluster of CL CTAs through distributed shared memory:
// every CTA writes its values into rank 0's table, rank 0 adds them in rank order.
// Returns true in (rank 0, thread 0), where out[0..NV) holds the totals.  CL == 1: no exchange.
template <int NV, int CL>
__device__ __forceinline__ bool cluster_sum(const double (&mine)[NV], double *table /* smem [CL][NV] */,
                                            double (&out)[NV]) {
  if (CL == 1) {
    if (threadIdx.x == 0) {
#pragma unroll
      for (int i = 0; i < NV; ++i) out[i] = mine[i];
      return true;
    }
    return false;
  }
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  // distributed shared memory of another CTA may only be touched once that CTA is known to
  // have started: all CTAs of the cluster meet here first
  cluster.sync();
  if (threadIdx.x == 0) {
    double *dst = cluster.map_shared_rank(table, 0);
#pragma unroll
    for (int i = 0; i < NV; ++i) dst[rank * NV + i] = mine[i];
  }
  cluster.sync();
  if (rank == 0 && threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      double a = 0.0;
      for (int r = 0; r < CL; ++r) a += table[r * NV + i];
      out[i] = a;
    }
    return true;
  }
  return false;
}

// CTAs per sample: eight (a thread-block cluster, <= 2 points per thread) while the batch is too
// small to fill the GPU -- the pass is then pure latency; two, then one (12 points per thread)
// as the batch grows and the per-CTA barriers / cluster exchange become the cost.
static inline int pm_cluster_for(int batch_size) {
  return batch_size >= 512 ? 1 : (batch_size >= 128 ? 2 : PM_CLUSTER);
}

// One cluster of CL blocks per sample.  est / loss_map / loss_sum are optional
// outputs; obs == nullptr turns the kernel into plain Transform3D forward.
template <int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(PM_THREADS)
pm_fwd_kernel(const float *__restrict__ points, const float *__restrict__ obs,
              const float *__restrict__ weights, const float *__restrict__ rotation,
              const float *__restrict__ trans, const float *__restrict__ pose_src,
              PoseParams pp, int loss_type, float norm_term, float sigma, float *est,
              float *__restrict__ loss_map, float *__restrict__ loss_sum, int N,
              int accumulate) {
  pdl_launch_dependents();
  const int b = blockIdx.x / CL;
  const int crank = blockIdx.x % CL;
  __shared__ PoseOut po;
  __shared__ double red[PM_THREADS / 32];
  __shared__ double table[CL];
  if (threadIdx.x == 0) compose_pose(rotation + b * 4, trans + b * 3, pose_src + b * 12, pp, po);
  __syncthreads();
  float R[9], T[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = po.R[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) T[i] = po.T[i];
  const float sigma2 = sigma * sigma, inv_sigma2 = 1.f / (sigma * sigma);
  const size_t base = (size_t)b * 3 * N;
  double acc = 0.0;
  for (int n = crank * PM_THREADS + threadIdx.x; n < N; n += PM_THREADS * CL) {
    const float p0 = __ldg(points + base + n);
    const float p1 = __ldg(points + base + N + n);
    const float p2 = __ldg(points + base + 2 * N + n);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      // batch_dot(Rm_tgt, X) + T_tgt, transform3d.py:113-114
      const float e = __fadd_rn(
          __fmaf_rn(R[k * 3 + 2], p2, __fmaf_rn(R[k * 3 + 1], p1, __fmul_rn(R[k * 3 + 0], p0))),
          T[k]);
      const size_t idx = base + (size_t)k * N + n;
      if (est) est[idx] = accumulate ? est[idx] + e : e;
      if (obs) {
        const float x = __fdiv_rn(__fsub_rn(e, __ldg(obs + idx)), norm_term);  // flownet.py:463
        const float l = __fmul_rn(__ldg(weights + idx), loss_value(x, loss_type, sigma2, inv_sigma2));
        if (loss_map) loss_map[idx] = l;
        acc += (double)l;
      }
    }
  }
  // every CTA of the cluster takes part in the reduction (cluster.sync inside)
  double part[1] = {acc}, mine[1], tot[1];
  block_sum_n<1>(part, red, mine);
  if (cluster_sum<1, CL>(mine, table, tot) && loss_sum) loss_sum[b] = (float)tot[0];
  // launched programmatically behind another grid and never reads it: wait HERE, after the
  // work, so that this grid's completion implies the predecessor's (a later normal launch
  // that waits for this one has then transitively waited for everything before it)
  pdl_wait();
}

// quat2mat_backward, transform3d.py:256-327 (float64 tail, float32 result).
__device__ void quat2mat_bwd(const double *D, const float *q, float *gq, int accumulate) {
  const float w = q[0], x = q[1], y = q[2], z = q[3];
  const float Nq = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w, w), __fmul_rn(x, x)), __fmul_rn(y, y)),
                             __fmul_rn(z, z));
  const float Nq_sqrt = sqrtf(Nq);
  const double dn = (double)Nq - 1.0;
  float out[4] = {0.f, 0.f, 0.f, 0.f};
  if (-1e-4 < dn && dn < 1e-4) {  // :263-264
    const double w_ = __fdiv_rn(w, Nq_sqrt), x_ = __fdiv_rn(x, Nq_sqrt),
                 y_ = __fdiv_rn(y, Nq_sqrt), z_ = __fdiv_rn(z, Nq_sqrt);
    const double s = 2.0;
    double wd = -z_ * D[1] + y_ * D[2] + z_ * D[3] - x_ * D[5] - y_ * D[6] + x_ * D[7];
    double xd = y_ * D[1] + z_ * D[2] + y_ * D[3] - 2 * x_ * D[4] - w_ * D[5] + z_ * D[6] +
                w_ * D[7] - 2 * x_ * D[8];
    double yd = -2 * y_ * D[0] + x_ * D[1] + w_ * D[2] + x_ * D[3] + z_ * D[5] - w_ * D[6] +
                z_ * D[7] - 2 * y_ * D[8];
    double zd = -2 * z_ * D[0] - w_ * D[1] + x_ * D[2] + w_ * D[3] - 2 * z_ * D[4] + y_ * D[5] +
                x_ * D[6] + y_ * D[7];
    wd *= s; xd *= s; yd *= s; zd *= s;
    const double ns = (double)Nq_sqrt;
    const double share = ns * ns * ns * (w * wd + x * xd + y * yd + z * zd);  // :316-318
    out[0] = (float)(ns * wd - w * share);
    out[1] = (float)(ns * xd - x * share);
    out[2] = (float)(ns * yd - y * share);
    out[3] = (float)(ns * zd - z * share);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) gq[i] = accumulate ? gq[i] + out[i] : out[i];
}

// One block per sample.  FROM_LOSS: the head gradient is computed from the
// loss definition (MakeLoss grad_scale, weights, d|x|/dx, 1/norm); otherwise
// it is read from out_grad (plain Transform3D backward).
template <bool FROM_LOSS, int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(PM_THREADS)
pm_bwd_kernel(const float *__restrict__ out_grad, const float *__restrict__ points,
              const float *__restrict__ obs, const float *__restrict__ weights,
              const float *__restrict__ rotation, const float *__restrict__ trans,
              const float *__restrict__ pose_src, PoseParams pp, int loss_type,
              float norm_term, float sigma, float grad_scale, float *__restrict__ grad_rot,
              float *__restrict__ grad_trans, int N, int accumulate) {
  pdl_launch_dependents();
  const int b = blockIdx.x / CL;
  const int crank = blockIdx.x % CL;
  __shared__ PoseOut po;
  __shared__ double red[(PM_THREADS / 32) * 12];
  __shared__ double table[CL * 12];
  if (threadIdx.x == 0) compose_pose(rotation + b * 4, trans + b * 3, pose_src + b * 12, pp, po);
  __syncthreads();
  float R[9], T[3], Rs[9], Ts[3];
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = po.R[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    T[i] = po.T[i];
    Rs[i * 3 + 0] = __ldg(pose_src + b * 12 + i * 4 + 0);
    Rs[i * 3 + 1] = __ldg(pose_src + b * 12 + i * 4 + 1);
    Rs[i * 3 + 2] = __ldg(pose_src + b * 12 + i * 4 + 2);
    Ts[i] = __ldg(pose_src + b * 12 + i * 4 + 3);
  }
  const float sigma2 = sigma * sigma, inv_sigma2 = 1.f / (sigma * sigma);
  const size_t base = (size_t)b * 3 * N;
  double gt[3] = {0, 0, 0};  // sum_n G            (transform3d.py:138)
  double gr[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // G X^T (:154-157)
  for (int n = crank * PM_THREADS + threadIdx.x; n < N; n += PM_THREADS * CL) {
    float p[3];
    p[0] = __ldg(points + base + n);
    p[1] = __ldg(points + base + N + n);
    p[2] = __ldg(points + base + 2 * N + n);
    float xr[3];  // right-hand factor of the R-gradient outer product
    if (pp.rot_coord == RC_ROT_NAIVE) {  // :166-173 uses Rm_src X + T_src
#pragma unroll
      for (int k = 0; k < 3; ++k)
        xr[k] = __fadd_rn(__fmaf_rn(Rs[k * 3 + 2], p[2],
                                    __fmaf_rn(Rs[k * 3 + 1], p[1], __fmul_rn(Rs[k * 3 + 0], p[0]))),
                          Ts[k]);
    } else {
      xr[0] = p[0]; xr[1] = p[1]; xr[2] = p[2];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const size_t idx = base + (size_t)k * N + n;
      float g;
      if (FROM_LOSS) {
        const float e = __fadd_rn(
            __fmaf_rn(R[k * 3 + 2], p[2], __fmaf_rn(R[k * 3 + 1], p[1], __fmul_rn(R[k * 3 + 0], p[0]))),
            T[k]);
        const float x = __fdiv_rn(__fsub_rn(e, __ldg(obs + idx)), norm_term);
        g = __fdiv_rn(__fmul_rn(__fmul_rn(grad_scale, __ldg(weights + idx)),
                                loss_grad(x, loss_type, sigma2, inv_sigma2)),
                      norm_term);
      } else {
        g = __ldg(out_grad + idx);
      }
      gt[k] += (double)g;
      gr[k * 3 + 0] += (double)g * xr[0];
      gr[k * 3 + 1] += (double)g * xr[1];
      gr[k * 3 + 2] += (double)g * xr[2];
    }
  }
  double part[12], mine[12], sums[12];
#pragma unroll
  for (int k = 0; k < 3; ++k) part[k] = gt[k];
#pragma unroll
  for (int k = 0; k < 9; ++k) part[3 + k] = gr[k];
  block_sum_n<12>(part, red, mine);  // one pass: twelve sums between two barriers
  const bool lead = cluster_sum<12, CL>(mine, table, sums);
  pdl_wait();  // see pm_fwd_kernel: this grid's completion must imply its predecessor's
  if (!lead) return;
  // values the reference holds as float32 NDArrays
  float Df[3], Gf[9];
#pragma unroll
  for (int k = 0; k < 3; ++k) Df[k] = (float)sums[k];
#pragma unroll
  for (int k = 0; k < 9; ++k) Gf[k] = (float)sums[3 + k];

  // ---- translation gradient, T_transform_backward :193-225
  float gtd[3];
  if (pp.rot_coord == RC_ROT_NAIVE) {
    gtd[0] = Df[0]; gtd[1] = Df[1]; gtd[2] = Df[2];  // :141-142
  } else {
    const float z2 = po.z2;
    if (pp.rot_coord == RC_ROT_CAMERA_NEW) {
      gtd[0] = Df[0] * (pp.stds[0] * Ts[2]);
      gtd[1] = Df[1] * (pp.stds[1] * Ts[2]);
      gtd[2] = Df[2] * (-pp.stds[2] * z2);
    } else {
      gtd[0] = Df[0] * (pp.stds[0] * z2);
      gtd[1] = Df[1] * (pp.stds[1] * z2);
      const float share = -pp.stds[2] * z2;
      gtd[2] = Df[0] * (share * (po.Td1[0] + Ts[0] / Ts[2])) +
               Df[1] * (share * (po.Td1[1] + Ts[1] / Ts[2])) + Df[2] * (-pp.stds[2] * z2);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k)
    grad_trans[b * 3 + k] = accumulate ? grad_trans[b * 3 + k] + gtd[k] : gtd[k];

  // ---- rotation gradient :154-184
  double Dq[9];
  if (pp.rot_coord == RC_ROT_MODEL) {  // Rm_src^T * G
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        Dq[i * 3 + j] = (double)(Rs[0 * 3 + i] * Gf[0 * 3 + j] + Rs[1 * 3 + i] * Gf[1 * 3 + j] +
                                 Rs[2 * 3 + i] * Gf[2 * 3 + j]);
  } else if (pp.rot_coord == RC_ROT_NAIVE) {
    for (int i = 0; i < 9; ++i) Dq[i] = (double)Gf[i];
  } else {  // camera / camera_new: G * Rm_src^T
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        Dq[i * 3 + j] = (double)(Gf[i * 3 + 0] * Rs[j * 3 + 0] + Gf[i * 3 + 1] * Rs[j * 3 + 1] +
                                 Gf[i * 3 + 2] * Rs[j * 3 + 2]);
  }
  quat2mat_bwd(Dq, rotation + b * 4, grad_rot + b * 4, accumulate);
}

static int fill_pose_params(PoseParams &pp, const float *means, const float *stds, int rot_coord) {
  if (rot_coord < RC_ROT_MODEL || rot_coord > RC_ROT_NAIVE) return RC_ERR_INVALID_ARG;
  for (int i = 0; i < 3; ++i) {
    pp.means[i] = means ? means[i] : 0.f;
    pp.stds[i] = stds ? stds[i] : 1.f;
  }
  pp.rot_coord = rot_coord;
  return RC_OK;
}

}  // namespace rc

// Launch KERNEL<..., CL> with CL = pm_cluster_for(batch) CTAs per sample.
#define RC_PM_DISPATCH(batch, LAUNCH)           \
  do {                                          \
    switch (rc::pm_cluster_for(batch)) {        \
      case 1: { constexpr int CL = 1; LAUNCH; } break; \
      case 2: { constexpr int CL = 2; LAUNCH; } break; \
      default: { constexpr int CL = rc::PM_CLUSTER; LAUNCH; } break; \
    }                                           \
  } while (0)

extern "C" {

int rc_transform3d_fwd_dev(const float *points, const float *rotation, const float *trans,
                           const float *pose_src, const float *T_means, const float *T_stds,
                           int rot_coord, float *out, int batch_size, int num_points,
                           int accumulate, void *stream) {
  RC_REQUIRE(batch_size >= 0 && num_points >= 0, RC_ERR_INVALID_ARG);
  rc::PoseParams pp;
  RC_REQUIRE(rc::fill_pose_params(pp, T_means, T_stds, rot_coord) == RC_OK, RC_ERR_INVALID_ARG);
  if (batch_size == 0 || num_points == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(points && rotation && trans && pose_src && out, RC_ERR_INVALID_ARG);
  RC_PM_DISPATCH(batch_size,
                 (rc::pm_fwd_kernel<CL><<<batch_size * CL, rc::PM_THREADS, 0,
                                          static_cast<cudaStream_t>(stream)>>>(
                     points, nullptr, nullptr, rotation, trans, pose_src, pp, RC_PM_L1, 1.f, 1.f,
                     out, nullptr, nullptr, num_points, accumulate)));
  return rc::note_status(rc::check_launch());
}

int rc_transform3d_bwd_dev(const float *out_grad, const float *points, const float *rotation,
                           const float *trans, const float *pose_src, const float *T_means,
                           const float *T_stds, int rot_coord, float *grad_rot,
                           float *grad_trans, int batch_size, int num_points, int accumulate,
                           void *stream) {
  RC_REQUIRE(batch_size >= 0 && num_points >= 0, RC_ERR_INVALID_ARG);
  rc::PoseParams pp;
  RC_REQUIRE(rc::fill_pose_params(pp, T_means, T_stds, rot_coord) == RC_OK, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(out_grad && points && rotation && trans && pose_src && grad_rot && grad_trans,
             RC_ERR_INVALID_ARG);
  RC_PM_DISPATCH(batch_size,
                 (rc::pm_bwd_kernel<false, CL><<<batch_size * CL, rc::PM_THREADS, 0,
                                                 static_cast<cudaStream_t>(stream)>>>(
                     out_grad, points, nullptr, nullptr, rotation, trans, pose_src, pp, RC_PM_L1,
                     1.f, 1.f, 1.f, grad_rot, grad_trans, num_points, accumulate)));
  return rc::note_status(rc::check_launch());
}

int rc_pm_loss_fwd_dev(const float *points, const float *points_obs, const float *weights,
                       const float *rotation, const float *trans, const float *pose_src,
                       const float *T_means, const float *T_stds, int rot_coord, int loss_type,
                       float norm_term, float sl1_scalar, float *est, float *loss_map,
                       float *loss_sum, int batch_size, int num_points, void *stream) {
  RC_REQUIRE(batch_size >= 0 && num_points >= 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(loss_type >= RC_PM_L1 && loss_type <= RC_PM_SMOOTH_L1, RC_ERR_INVALID_ARG);
  RC_REQUIRE(norm_term != 0.f && sl1_scalar != 0.f, RC_ERR_INVALID_ARG);
  rc::PoseParams pp;
  RC_REQUIRE(rc::fill_pose_params(pp, T_means, T_stds, rot_coord) == RC_OK, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(points && points_obs && weights && rotation && trans && pose_src, RC_ERR_INVALID_ARG);
  // rc::tl_pdl: the fused step issues the forward beside the backward (neither reads the other)
  RC_PM_DISPATCH(batch_size,
                 RC_CUDA_TRY(rc::launch_maybe_pdl(
                     rc::pm_fwd_kernel<CL>, dim3(batch_size * CL), dim3(rc::PM_THREADS), 0,
                     static_cast<cudaStream_t>(stream), rc::tl_pdl, points, points_obs, weights,
                     rotation, trans, pose_src, pp, loss_type, norm_term, sl1_scalar, est,
                     loss_map, loss_sum, num_points, 0)));
  return rc::note_status(rc::check_launch());
}

int rc_pm_loss_bwd_dev(const float *points, const float *points_obs, const float *weights,
                       const float *rotation, const float *trans, const float *pose_src,
                       const float *T_means, const float *T_stds, int rot_coord, int loss_type,
                       float norm_term, float sl1_scalar, float grad_scale, float *grad_rot,
                       float *grad_trans, int batch_size, int num_points, void *stream) {
  RC_REQUIRE(batch_size >= 0 && num_points >= 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(loss_type >= RC_PM_L1 && loss_type <= RC_PM_SMOOTH_L1, RC_ERR_INVALID_ARG);
  RC_REQUIRE(norm_term != 0.f && sl1_scalar != 0.f, RC_ERR_INVALID_ARG);
  rc::PoseParams pp;
  RC_REQUIRE(rc::fill_pose_params(pp, T_means, T_stds, rot_coord) == RC_OK, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(points && points_obs && weights && rotation && trans && pose_src && grad_rot &&
                 grad_trans,
             RC_ERR_INVALID_ARG);
  RC_PM_DISPATCH(batch_size,
                 (rc::pm_bwd_kernel<true, CL><<<batch_size * CL, rc::PM_THREADS, 0,
                                                static_cast<cudaStream_t>(stream)>>>(
                     nullptr, points, points_obs, weights, rotation, trans, pose_src, pp,
                     loss_type, norm_term, sl1_scalar, grad_scale, grad_rot, grad_trans,
                     num_points, 0)));
  return rc::note_status(rc::check_launch());
}

}  // extern "C"
