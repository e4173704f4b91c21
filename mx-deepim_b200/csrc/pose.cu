// pose.cu -- per-pair pose algebra between refinement iterations, batched on device.
//
// Replaces the per-pair Python loop of
//   /root/reference/lib/pair_matching/batch_updater_py_multi.py:211-312
// i.e. RT_transform (lib/pair_matching/RT_transform.py:135-161), calc_RT_delta (:16-48,
// mat2quat :446-523), calc_se3 (:186-197 -> lib/utils/projection.py:12-43) and
// KT = K . [R_rel | t_rel] (batch_updater_py_multi.py:306-312).
//
// B x a few hundred scalar operations: one thread per pair, float64 where the reference's
// numpy evaluates in float64, float32 where it stores float32 (se3_inverse / se3_mul outputs,
// the MXNet blobs).  Keeping it on the device removes the .asnumpy() round trips
// (:144-174) that serialise the reference's loop.
#include "common.cuh"

namespace rc {

// Largest-eigenvalue eigenvector of a symmetric 4x4 matrix by cyclic Jacobi rotations.
__device__ void sym4_max_eigvec(double A[4][4], double v[4]) {
  double V[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
  for (int sweep = 0; sweep < 24; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 4; ++p)
      for (int q = p + 1; q < 4; ++q) off += A[p][q] * A[p][q];
    if (off < 1e-30) break;
    for (int p = 0; p < 4; ++p) {
      for (int q = p + 1; q < 4; ++q) {
        if (fabs(A[p][q]) < 1e-300) continue;
        const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 4; ++k) {
          const double akp = A[k][p], akq = A[k][q];
          A[k][p] = c * akp - s * akq;
          A[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 4; ++k) {
          const double apk = A[p][k], aqk = A[q][k];
          A[p][k] = c * apk - s * aqk;
          A[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 4; ++k) {
          const double vkp = V[k][p], vkq = V[k][q];
          V[k][p] = c * vkp - s * vkq;
          V[k][q] = s * vkp + c * vkq;
        }
      }
    }
  }
  int best = 0;
  for (int i = 1; i < 4; ++i)
    if (A[i][i] > A[best][best]) best = i;
  for (int k = 0; k < 4; ++k) v[k] = V[k][best];
}

// Fast path for the matrices mat2quat actually sees.  For a rotation matrix the 4x4 K of
// RT_transform.py:497-507 has the eigenvalues {1, -1/3, -1/3, -1/3}; the float32-rounded poses
// of the path perturb them by ~1e-7.  Power iteration on K + I/3 therefore contracts everything
// but the wanted eigenvector by ~1e-7 per step: five steps from the largest column reach double
// precision, where 24 Jacobi sweeps are a ~35k-cycle dependent chain on ONE thread per pair (the
// whole refine iteration waits for this kernel).  The result is accepted only if it is an
// eigenvector to 1e-14 AND its Rayleigh quotient dominates the spectrum bound of the rest
// (trace argument); anything else (a far-from-orthonormal M) falls back to Jacobi.
__device__ bool sym4_max_eigvec_shifted_power(const double K[4][4], double v[4]) {
  double S[4][4];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) S[i][j] = K[i][j] + (i == j ? 1.0 / 3.0 : 0.0);
  int best = 0;
  double bn = -1.0;
  for (int j = 0; j < 4; ++j) {
    const double n = S[0][j] * S[0][j] + S[1][j] * S[1][j] + S[2][j] * S[2][j] + S[3][j] * S[3][j];
    if (n > bn) {
      bn = n;
      best = j;
    }
  }
  if (!(bn > 1e-4)) return false;
  double x[4] = {S[0][best], S[1][best], S[2][best], S[3][best]};
  for (int it = 0; it < 5; ++it) {
    double y[4];
    for (int i = 0; i < 4; ++i) y[i] = S[i][0] * x[0] + S[i][1] * x[1] + S[i][2] * x[2] + S[i][3] * x[3];
    const double inv = rsqrt(y[0] * y[0] + y[1] * y[1] + y[2] * y[2] + y[3] * y[3]);
    for (int i = 0; i < 4; ++i) x[i] = y[i] * inv;
  }
  // residual of K x = lambda x
  double kx[4], lam = 0.0;
  for (int i = 0; i < 4; ++i) {
    kx[i] = K[i][0] * x[0] + K[i][1] * x[1] + K[i][2] * x[2] + K[i][3] * x[3];
    lam += kx[i] * x[i];
  }
  double res = 0.0;
  for (int i = 0; i < 4; ++i) res += (kx[i] - lam * x[i]) * (kx[i] - lam * x[i]);
  // the other three eigenvalues sum to trace - lam; with sum of squares ||K||_F^2 - lam^2 none of
  // them can exceed sqrt(that): lam must beat it to be the largest
  double fro = 0.0;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) fro += K[i][j] * K[i][j];
  const double rest = fro - lam * lam;
  if (!(res < 1e-28) || !(lam > 0.0) || !(rest < lam * lam)) return false;
  for (int i = 0; i < 4; ++i) v[i] = x[i];
  return true;
}

// mat2quat, RT_transform.py:446-523 (w, x, y, z with w >= 0).
__device__ void mat2quat(const double M[9], double q[4]) {
  const double Qxx = M[0], Qyx = M[1], Qzx = M[2], Qxy = M[3], Qyy = M[4], Qzy = M[5], Qxz = M[6],
               Qyz = M[7], Qzz = M[8];
  double K[4][4];
  K[0][0] = (Qxx - Qyy - Qzz) / 3.0;
  K[1][0] = (Qyx + Qxy) / 3.0;
  K[1][1] = (Qyy - Qxx - Qzz) / 3.0;
  K[2][0] = (Qzx + Qxz) / 3.0;
  K[2][1] = (Qzy + Qyz) / 3.0;
  K[2][2] = (Qzz - Qxx - Qyy) / 3.0;
  K[3][0] = (Qyz - Qzy) / 3.0;
  K[3][1] = (Qzx - Qxz) / 3.0;
  K[3][2] = (Qxy - Qyx) / 3.0;
  K[3][3] = (Qxx + Qyy + Qzz) / 3.0;
  for (int i = 0; i < 4; ++i)
    for (int j = i + 1; j < 4; ++j) K[i][j] = K[j][i];  // eigh reads the lower triangle
  double v[4];
  if (!sym4_max_eigvec_shifted_power(K, v)) sym4_max_eigvec(K, v);
  q[0] = v[3];
  q[1] = v[0];
  q[2] = v[1];
  q[3] = v[2];
  if (q[0] < 0.0)
    for (int i = 0; i < 4; ++i) q[i] = -q[i];
}

struct PoseUpdateParams {
  float means[3], stds[3];
  int rot_coord;
};

// Stand-in for the FlowNet heads in synthetic refine loops ("predicted (q, t) =
// calc_RT_delta(src, tgt) damped by 0.5 + N(0, small)").  The CNN is out of scope; the stand-in
// must not weigh on the measured loop, so it either runs as one small launch
// (rc_standin_predict_dev) or rides at the head of the pose update (rc_pose_update_standin_dev),
// which takes it off the loop's chain of dependent launches altogether.
struct Standin {
  const float *rot_gt, *trans_gt, *noise_q, *noise_t;  // rot_gt == NULL: no stand-in
  float *rot_est, *trans_est;                          // written for the loss (and the tests)
  float damping;
};

__device__ __forceinline__ void standin_predict(const Standin &sp, int b, float *q, float *t) {
  float n2 = 0.f;
  for (int i = 0; i < 4; ++i) {
    q[i] = __fadd_rn(__fmaf_rn(sp.damping, sp.rot_gt[b * 4 + i],
                               __fmul_rn(1.f - sp.damping, i == 0 ? 1.f : 0.f)),
                     sp.noise_q[b * 4 + i]);
    n2 = __fmaf_rn(q[i], q[i], n2);
  }
  const float inv = __fdiv_rn(1.f, sqrtf(n2));  // L2Normalization upstream of the pose update
  for (int i = 0; i < 4; ++i) q[i] = __fmul_rn(q[i], inv);
  for (int i = 0; i < 3; ++i)
    t[i] = __fmaf_rn(sp.damping, sp.trans_gt[b * 3 + i], sp.noise_t[b * 3 + i]);
}

__global__ void pose_update_kernel(const float *__restrict__ src_pose, const float *rot_est,
                                   const float *trans_est,
                                   const float *__restrict__ tgt_pose, const float *__restrict__ K,
                                   PoseUpdateParams pp, Standin sp, float *__restrict__ refined_pose,
                                   float *rot_res, float *trans_res,
                                   float *__restrict__ KT, int B, int parts) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  // ---- RT_transform (RT_transform.py:135-161): quat = r / ||r|| in float32, quat2mat float64
  float r[4], te[3];
  if (sp.rot_gt) {  // the stand-in's prediction from the current labels (may alias rot_res)
    standin_predict(sp, b, r, te);
    for (int i = 0; i < 4; ++i) sp.rot_est[b * 4 + i] = r[i];
    for (int i = 0; i < 3; ++i) sp.trans_est[b * 3 + i] = te[i];
  } else {
    for (int i = 0; i < 4; ++i) r[i] = rot_est[b * 4 + i];
    for (int i = 0; i < 3; ++i) te[i] = trans_est[b * 3 + i];
  }
  const float nr = sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r[0], r[0]), __fmul_rn(r[1], r[1])),
                                             __fmul_rn(r[2], r[2])),
                                   __fmul_rn(r[3], r[3])));
  float qf[4];
  for (int i = 0; i < 4; ++i) qf[i] = __fdiv_rn(r[i], nr);
  const float Nqf = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(qf[0], qf[0]), __fmul_rn(qf[1], qf[1])),
                                        __fmul_rn(qf[2], qf[2])),
                              __fmul_rn(qf[3], qf[3]));
  double Rd[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if ((double)Nqf >= 2.220446049250313e-16) {  // _FLOAT_EPS, :247,:425
    const double w = qf[0], x = qf[1], y = qf[2], z = qf[3];
    const double s = 2.0 / (double)Nqf;
    const double X = x * s, Y = y * s, Z = z * s;
    Rd[0] = 1.0 - (y * Y + z * Z); Rd[1] = x * Y - w * Z;         Rd[2] = x * Z + w * Y;
    Rd[3] = x * Y + w * Z;         Rd[4] = 1.0 - (x * X + z * Z); Rd[5] = y * Z - w * X;
    Rd[6] = x * Z - w * Y;         Rd[7] = y * Z + w * X;         Rd[8] = 1.0 - (x * X + y * Y);
  }
  double Rs[9], Ts[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) Rs[i * 3 + j] = src_pose[b * 12 + i * 4 + j];
    Ts[i] = src_pose[b * 12 + i * 4 + 3];
  }
  double Rn[9];  // R_transform, :51-69
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double a = 0.0;
      for (int k = 0; k < 3; ++k)
        a += (pp.rot_coord == RC_ROT_MODEL) ? Rs[i * 3 + k] * Rd[k * 3 + j] : Rd[i * 3 + k] * Rs[k * 3 + j];
      Rn[i * 3 + j] = a;
    }
  // T_transform, :82-103, float32 arithmetic on float32 operands
  float td1[3];
  for (int i = 0; i < 3; ++i)
    td1[i] = __fadd_rn(__fmul_rn(te[i], pp.stds[i]), pp.means[i]);
  const float Tsf[3] = {(float)Ts[0], (float)Ts[1], (float)Ts[2]};
  const float z2 = __fdiv_rn(Tsf[2], expf(td1[2]));
  float Tn[3];
  Tn[2] = z2;
  if (pp.rot_coord == RC_ROT_CAMERA_NEW) {
    Tn[0] = __fadd_rn(__fmul_rn(Tsf[2], td1[0]), Tsf[0]);
    Tn[1] = __fadd_rn(__fmul_rn(Tsf[2], td1[1]), Tsf[1]);
  } else {
    Tn[0] = __fmul_rn(z2, __fadd_rn(td1[0], __fdiv_rn(Tsf[0], Tsf[2])));
    Tn[1] = __fmul_rn(z2, __fadd_rn(td1[1], __fdiv_rn(Tsf[1], Tsf[2])));
  }
  // the refined pose is stored into a float32 blob (update_data_batch, :389-397)
  float Pf[12];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) Pf[i * 4 + j] = (float)Rn[i * 3 + j];
    Pf[i * 4 + 3] = Tn[i];
  }
  // parts: RC_POSE_PART_POSE = the refined pose (all the rasteriser waits for),
  // RC_POSE_PART_LABELS = the new regression labels and KT (the float64 eigen-solve of mat2quat
  // is most of this kernel's latency; a caller may run it beside the render)
  if (parts & RC_POSE_PART_POSE)
    for (int i = 0; i < 12; ++i) refined_pose[b * 12 + i] = Pf[i];
  if (!(parts & RC_POSE_PART_LABELS)) return;

  // ---- calc_RT_delta(refined_pose (float64), tgt_pose) -> new regression labels (:288-295)
  double Rt[9], Tt[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) Rt[i * 3 + j] = tgt_pose[b * 12 + i * 4 + j];
    Tt[i] = tgt_pose[b * 12 + i * 4 + 3];
  }
  const double Tr[3] = {(double)Tn[0], (double)Tn[1], (double)Tn[2]};
  double Rdel[9];  // R_inv_transform, :72-79
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double a = 0.0;
      for (int k = 0; k < 3; ++k)
        a += (pp.rot_coord == RC_ROT_MODEL) ? Rn[k * 3 + i] * Rt[k * 3 + j] : Rt[i * 3 + k] * Rn[j * 3 + k];
      Rdel[i * 3 + j] = a;
    }
  double q[4];
  mat2quat(Rdel, q);
  double tdl[3];  // T_inv_transform, :113-132
  if (pp.rot_coord == RC_ROT_CAMERA_NEW) {
    tdl[0] = (Tt[0] - Tr[0]) / Tr[2];
    tdl[1] = (Tt[1] - Tr[1]) / Tr[2];
  } else {
    tdl[0] = Tt[0] / Tt[2] - Tr[0] / Tr[2];
    tdl[1] = Tt[1] / Tt[2] - Tr[1] / Tr[2];
  }
  tdl[2] = log(Tr[2] / Tt[2]);
  for (int i = 0; i < 4; ++i) rot_res[b * 4 + i] = (float)q[i];
  for (int i = 0; i < 3; ++i)
    trans_res[b * 3 + i] = (float)((tdl[i] - (double)pp.means[i]) / (double)pp.stds[i]);

  // ---- calc_se3 + KT (:306-312): se3_inverse and se3_mul round to float32 (projection.py:20,39)
  float inv[12];
  for (int i = 0; i < 3; ++i) {
    double a = 0.0;
    for (int j = 0; j < 3; ++j) {
      inv[i * 4 + j] = (float)Rn[j * 3 + i];
      a += Rn[j * 3 + i] * Tr[j];
    }
    inv[i * 4 + 3] = (float)(-1.0 * a);
  }
  float rel[12];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) {
      float a = 0.f;  // float32 x float32 matmul
      for (int k = 0; k < 3; ++k) a = __fmaf_rn((float)Rt[i * 3 + k], inv[k * 4 + j], a);
      rel[i * 4 + j] = a;
    }
    float a = 0.f;
    for (int k = 0; k < 3; ++k) a = __fmaf_rn((float)Rt[i * 3 + k], inv[k * 4 + 3], a);
    rel[i * 4 + 3] = __fadd_rn(a, (float)Tt[i]);
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 4; ++j) {
      double a = 0.0;
      for (int k = 0; k < 3; ++k) a += (double)K[i * 3 + k] * (double)rel[k * 4 + j];
      KT[b * 12 + i * 4 + j] = (float)a;  // .astype(np.float32) at :343
    }
}

// calc_RT_delta(pose_src, pose_tgt, T_means, T_stds, rot_coord, "QUAT") alone
// (RT_transform.py:16-48): the regression labels of ITERATION 0, which the reference's loader
// computes on the host per pair (lib/dataset/*: calc_RT_delta on the stored float32 poses).
__global__ void pose_delta_kernel(const float *__restrict__ src_pose,
                                  const float *__restrict__ tgt_pose, PoseUpdateParams pp,
                                  float *__restrict__ rot_res, float *__restrict__ trans_res, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double Rn[9], Tr[3], Rt[9], Tt[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) {
      Rn[i * 3 + j] = src_pose[b * 12 + i * 4 + j];
      Rt[i * 3 + j] = tgt_pose[b * 12 + i * 4 + j];
    }
    Tr[i] = src_pose[b * 12 + i * 4 + 3];
    Tt[i] = tgt_pose[b * 12 + i * 4 + 3];
  }
  double Rdel[9];  // R_inv_transform, :72-79
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double a = 0.0;
      for (int k = 0; k < 3; ++k)
        a += (pp.rot_coord == RC_ROT_MODEL) ? Rn[k * 3 + i] * Rt[k * 3 + j] : Rt[i * 3 + k] * Rn[j * 3 + k];
      Rdel[i * 3 + j] = a;
    }
  double q[4];
  mat2quat(Rdel, q);
  double tdl[3];  // T_inv_transform, :113-132
  if (pp.rot_coord == RC_ROT_CAMERA_NEW) {
    tdl[0] = (Tt[0] - Tr[0]) / Tr[2];
    tdl[1] = (Tt[1] - Tr[1]) / Tr[2];
  } else {
    tdl[0] = Tt[0] / Tt[2] - Tr[0] / Tr[2];
    tdl[1] = Tt[1] / Tt[2] - Tr[1] / Tr[2];
  }
  tdl[2] = log(Tr[2] / Tt[2]);
  for (int i = 0; i < 4; ++i) rot_res[b * 4 + i] = (float)q[i];
  for (int i = 0; i < 3; ++i)
    trans_res[b * 3 + i] = (float)((tdl[i] - (double)pp.means[i]) / (double)pp.stds[i]);
}

__global__ void standin_predict_kernel(Standin sp, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float q[4], t[3];
  standin_predict(sp, b, q, t);
  for (int i = 0; i < 4; ++i) sp.rot_est[b * 4 + i] = q[i];
  for (int i = 0; i < 3; ++i) sp.trans_est[b * 3 + i] = t[i];
}

}  // namespace rc

extern "C" {

int rc_pose_update_dev(const float *src_pose, const float *rot_est, const float *trans_est,
                       const float *tgt_pose, const float *K, const float *T_means,
                       const float *T_stds, int rot_coord, float *refined_pose, float *rot_res,
                       float *trans_res, float *KT, int batch_size, void *stream) {
  RC_REQUIRE(batch_size >= 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(rot_coord == RC_ROT_MODEL || rot_coord == RC_ROT_CAMERA || rot_coord == RC_ROT_CAMERA_NEW,
             RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(src_pose && rot_est && trans_est && tgt_pose && K && refined_pose && rot_res &&
                 trans_res && KT,
             RC_ERR_INVALID_ARG);
  rc::PoseUpdateParams pp;
  for (int i = 0; i < 3; ++i) {
    pp.means[i] = T_means ? T_means[i] : 0.f;
    pp.stds[i] = T_stds ? T_stds[i] : 1.f;
  }
  pp.rot_coord = rot_coord;
  rc::Standin none = {};
  rc::pose_update_kernel<<<(batch_size + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      src_pose, rot_est, trans_est, tgt_pose, K, pp, none, refined_pose, rot_res, trans_res, KT,
      batch_size, RC_POSE_PART_POSE | RC_POSE_PART_LABELS);
  return rc::note_status(rc::check_launch());
}

int rc_pose_update_standin_dev(const float *src_pose, const float *rot_gt, const float *trans_gt,
                               const float *noise_q, const float *noise_t, float damping,
                               float *rot_est, float *trans_est, const float *tgt_pose,
                               const float *K, const float *T_means, const float *T_stds,
                               int rot_coord, float *refined_pose, float *rot_res,
                               float *trans_res, float *KT, int parts, int batch_size,
                               void *stream) {
  RC_REQUIRE(batch_size >= 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(rot_coord == RC_ROT_MODEL || rot_coord == RC_ROT_CAMERA || rot_coord == RC_ROT_CAMERA_NEW,
             RC_ERR_INVALID_ARG);
  RC_REQUIRE(parts >= 1 && parts <= 3, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  // rot_gt == NULL: no stand-in, rot_est / trans_est are the update as given
  RC_REQUIRE(src_pose && rot_est && trans_est && tgt_pose && K, RC_ERR_INVALID_ARG);
  RC_REQUIRE(rot_gt == nullptr || (trans_gt && noise_q && noise_t), RC_ERR_INVALID_ARG);
  RC_REQUIRE(!(parts & RC_POSE_PART_POSE) || refined_pose, RC_ERR_INVALID_ARG);
  RC_REQUIRE(!(parts & RC_POSE_PART_LABELS) || (rot_res && trans_res && KT), RC_ERR_INVALID_ARG);
  rc::PoseUpdateParams pp;
  for (int i = 0; i < 3; ++i) {
    pp.means[i] = T_means ? T_means[i] : 0.f;
    pp.stds[i] = T_stds ? T_stds[i] : 1.f;
  }
  pp.rot_coord = rot_coord;
  rc::Standin sp = {rot_gt, trans_gt, noise_q, noise_t, rot_est, trans_est, damping};
  rc::pose_update_kernel<<<(batch_size + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      src_pose, rot_est, trans_est, tgt_pose, K, pp, sp, refined_pose, rot_res, trans_res, KT,
      batch_size, parts);
  return rc::note_status(rc::check_launch());
}

int rc_pose_delta_dev(const float *src_pose, const float *tgt_pose, const float *T_means,
                      const float *T_stds, int rot_coord, float *rot_res, float *trans_res,
                      int batch_size, void *stream) {
  RC_REQUIRE(batch_size >= 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(rot_coord == RC_ROT_MODEL || rot_coord == RC_ROT_CAMERA || rot_coord == RC_ROT_CAMERA_NEW,
             RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(src_pose && tgt_pose && rot_res && trans_res, RC_ERR_INVALID_ARG);
  rc::PoseUpdateParams pp;
  for (int i = 0; i < 3; ++i) {
    pp.means[i] = T_means ? T_means[i] : 0.f;
    pp.stds[i] = T_stds ? T_stds[i] : 1.f;
  }
  pp.rot_coord = rot_coord;
  rc::pose_delta_kernel<<<(batch_size + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      src_pose, tgt_pose, pp, rot_res, trans_res, batch_size);
  return rc::note_status(rc::check_launch());
}

int rc_standin_predict_dev(const float *rot_gt, const float *trans_gt, const float *noise_q,
                           const float *noise_t, float damping, float *rot_est, float *trans_est,
                           int batch_size, void *stream) {
  RC_REQUIRE(batch_size >= 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(rot_gt && trans_gt && noise_q && noise_t && rot_est && trans_est, RC_ERR_INVALID_ARG);
  rc::Standin sp = {rot_gt, trans_gt, noise_q, noise_t, rot_est, trans_est, damping};
  rc::standin_predict_kernel<<<(batch_size + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      sp, batch_size);
  return rc::note_status(rc::check_launch());
}

}  // extern "C"
