// Batched pose Kalman filter: the per-sequence loop of src/helper/postprocess.py:35-61 over
// the reference's Kalman_Filter (src/estimation/filter.py:107-202) with its State_R3xQuat pose algebra
// (src/estimation/state.py:100-185) and Linear_Motion model (src/estimation/motion.py:15-60), for S independent
// sequences at once: one thread per sequence, float64 like the reference (numpy), the T steps in order.
//
// State = t xyz | q xyzw (7).  Per observation:  update (K = P (P + R)^-1 with H = I, state += K (obs - state),
// quaternion re-normalised, covariance updated ELEMENT-WISE as the reference does, filter.py:186, which leaves it
// diagonal) -> predict (motion = relative transform of the last two states scaled by a_t / slerped by a_r, state
// = last o motion, covariance = A P A^T + Q).
// Quaternion <-> matrix conversions follow scipy's Rotation exactly (from_quat normalises; from_matrix branches on
// the largest of m00, m11, m22, trace and does NOT canonicalise the sign) because the filter does linear arithmetic on
// quaternion components.  ONE stated deviation: A is the analytic Jacobian of the composition; the reference
// differences it with h = machine epsilon, which yields rounding noise (profiles/r2_kalman_reference_probe.txt).
//
// Bound: latency (a few thousand dependent FP64 operations per step and sequence, 7x7 matrices in local memory);
// throughput comes from the number of sequences in flight.
#include "common.cuh"
#include <math_constants.h>

namespace plr2 {

namespace kf {

__device__ __forceinline__ void quat_to_matrix(const double *q, double *m) {
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double x = q[0] / n, y = q[1] / n, z = q[2] / n, w = q[3] / n;
  const double x2 = x * x, y2 = y * y, z2 = z * z, w2 = w * w;
  const double xy = x * y, zw = z * w, xz = x * z, yw = y * w, yz = y * z, xw = x * w;
  m[0] = x2 - y2 - z2 + w2; m[1] = 2 * (xy - zw);       m[2] = 2 * (xz + yw);
  m[3] = 2 * (xy + zw);      m[4] = -x2 + y2 - z2 + w2; m[5] = 2 * (yz - xw);
  m[6] = 2 * (xz - yw);      m[7] = 2 * (yz + xw);      m[8] = -x2 - y2 + z2 + w2;
}

__device__ __forceinline__ void matrix_to_quat(const double *m, double *q) {
  const double tr = m[0] + m[4] + m[8];
  const double dec[4] = {m[0], m[4], m[8], tr};
  int c = 0;                                   // np.argmax: first index of the maximum
#pragma unroll
  for (int i = 1; i < 4; ++i)
    if (dec[i] > dec[c]) c = i;
  if (c != 3) {
    const int i = c, j = (c + 1) % 3, k = (c + 2) % 3;
    q[i] = 1 - tr + 2 * m[3 * i + i];
    q[j] = m[3 * j + i] + m[3 * i + j];
    q[k] = m[3 * k + i] + m[3 * i + k];
    q[3] = m[3 * k + j] - m[3 * j + k];
  } else {
    q[0] = m[7] - m[5];
    q[1] = m[2] - m[6];
    q[2] = m[3] - m[1];
    q[3] = 1 + tr;
  }
  const double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
  for (int i = 0; i < 4; ++i) q[i] /= n;
}

__device__ __forceinline__ void matmul3(const double *a, const double *b, double *c) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) c[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
}

// out = SE3(p1) @ SE3(p2) as a state (state.py:137-139)
__device__ void apply_transform(const double *p1, const double *p2, double *out) {
  double r1[9], r2[9], r[9];
  quat_to_matrix(p1 + 3, r1);
  quat_to_matrix(p2 + 3, r2);
#pragma unroll
  for (int i = 0; i < 3; ++i) out[i] = r1[3 * i] * p2[0] + r1[3 * i + 1] * p2[1] + r1[3 * i + 2] * p2[2] + p1[i];
  matmul3(r1, r2, r);
  matrix_to_quat(r, out + 3);
}

// out = inv(p1) o p2 (state.py:141-147)
__device__ void relative_transform(const double *p1, const double *p2, double *out) {
  double r[9], rt[9], inv[7];
  quat_to_matrix(p1 + 3, r);
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) rt[3 * i + j] = r[3 * j + i];
#pragma unroll
  for (int i = 0; i < 3; ++i) inv[i] = -(rt[3 * i] * p1[0] + rt[3 * i + 1] * p1[1] + rt[3 * i + 2] * p1[2]);
  matrix_to_quat(rt, inv + 3);
  apply_transform(inv, p2, out);
}

// d apply_transform(p1, p2) / d p1, analytic: t = t1 + R(q1/|q1|) t2, q = s (q1/|q1|) (x) (q2/|q2|)
__device__ void jacobian(const double *p1, const double *p2, const double *q_out, double (*J)[7]) {
  const double n1 = sqrt(p1[3] * p1[3] + p1[4] * p1[4] + p1[5] * p1[5] + p1[6] * p1[6]);
  const double u[4] = {p1[3] / n1, p1[4] / n1, p1[5] / n1, p1[6] / n1};
  double proj[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) proj[i][j] = ((i == j ? 1.0 : 0.0) - u[i] * u[j]) / n1;
  const double x = u[0], y = u[1], z = u[2], w = u[3], a = p2[0], b = p2[1], c = p2[2];
  const double dR[3][4] = {
      {2 * (x * a + y * b + z * c), 2 * (-y * a + x * b + w * c), 2 * (-z * a - w * b + x * c), 2 * (w * a - z * b + y * c)},
      {2 * (y * a - x * b - w * c), 2 * (x * a + y * b + z * c), 2 * (w * a - z * b + y * c), 2 * (z * a + w * b - x * c)},
      {2 * (z * a + w * b - x * c), 2 * (-w * a + z * b - y * c), 2 * (x * a + y * b + z * c), 2 * (-y * a + x * b + w * c)}};
  const double n2 = sqrt(p2[3] * p2[3] + p2[4] * p2[4] + p2[5] * p2[5] + p2[6] * p2[6]);
  const double bx = p2[3] / n2, by = p2[4] / n2, bz = p2[5] / n2, bw = p2[6] / n2;
  const double right[4][4] = {{bw, bz, -by, bx}, {-bz, bw, bx, by}, {by, -bx, bw, bz}, {-bx, -by, -bz, bw}};
  double dot = 0.0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double pr = 0.0;
#pragma unroll
    for (int j = 0; j < 4; ++j) pr += right[i][j] * u[j];
    dot += q_out[i] * pr;
  }
  const double s = dot >= 0 ? 1.0 : -1.0;
#pragma unroll
  for (int i = 0; i < 7; ++i)
#pragma unroll
    for (int j = 0; j < 7; ++j) J[i][j] = 0.0;
#pragma unroll
  for (int i = 0; i < 3; ++i) J[i][i] = 1.0;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double v = 0.0;
#pragma unroll
      for (int k = 0; k < 4; ++k) v += dR[i][k] * proj[k][j];
      J[i][3 + j] = v;
    }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double v = 0.0;
#pragma unroll
      for (int k = 0; k < 4; ++k) v += right[i][k] * proj[k][j];
      J[3 + i][3 + j] = s * v;
    }
}

// inverse of a 7x7 matrix by Gauss-Jordan elimination with partial pivoting; returns false if singular
__device__ bool invert7(double (*a)[7], double (*inv)[7]) {
  for (int i = 0; i < 7; ++i)
    for (int j = 0; j < 7; ++j) inv[i][j] = i == j ? 1.0 : 0.0;
  for (int c = 0; c < 7; ++c) {
    int p = c;
    double best = fabs(a[c][c]);
    for (int r = c + 1; r < 7; ++r)
      if (fabs(a[r][c]) > best) { best = fabs(a[r][c]); p = r; }
    if (!(best > 0.0)) return false;
    if (p != c)
      for (int j = 0; j < 7; ++j) {
        double t = a[c][j]; a[c][j] = a[p][j]; a[p][j] = t;
        t = inv[c][j]; inv[c][j] = inv[p][j]; inv[p][j] = t;
      }
    const double d = 1.0 / a[c][c];
    for (int j = 0; j < 7; ++j) { a[c][j] *= d; inv[c][j] *= d; }
    for (int r = 0; r < 7; ++r) {
      if (r == c) continue;
      const double f = a[r][c];
      if (f == 0.0) continue;
      for (int j = 0; j < 7; ++j) { a[r][j] -= f * a[c][j]; inv[r][j] -= f * inv[c][j]; }
    }
  }
  return true;
}

}  // namespace kf

__global__ void __launch_bounds__(64)
kalman_filter_kernel(const double *__restrict__ prior, const double *__restrict__ prior_var,
                     const double *__restrict__ obs, const double *__restrict__ sensor_var,
                     const double *__restrict__ q_diag, const int32_t *__restrict__ lengths, double a_t, double a_r,
                     int S, int T, double *__restrict__ filtered, double *__restrict__ p_diag,
                     double *__restrict__ predicted) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= S) return;
  int len = lengths ? lengths[s] : T;
  len = len < 0 ? 0 : (len > T ? T : len);
  double prev[7], cur[7];                     // the last two entries of the reference's `predictions` list
  bool have_prev = false;
  double P[7][7], Sm[7][7], Si[7][7];
#pragma unroll
  for (int i = 0; i < 7; ++i) cur[i] = prior[(size_t)s * 7 + i];
  for (int i = 0; i < 7; ++i)
    for (int j = 0; j < 7; ++j) P[i][j] = i == j ? prior_var[(size_t)s * 7 + i] : 0.0;
  double Q[7];
#pragma unroll
  for (int i = 0; i < 7; ++i) Q[i] = q_diag[i];
  for (int t = 0; t < len; ++t) {
    const size_t o = ((size_t)s * T + t) * 7;
    // ---- update (filter.py:159-189)
    for (int i = 0; i < 7; ++i)
      for (int j = 0; j < 7; ++j) Sm[i][j] = P[i][j] + (i == j ? sensor_var[o + i] : 0.0);
    const bool ok = kf::invert7(Sm, Si);
    double res[7], pd[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) res[i] = obs[o + i] - cur[i];
    for (int i = 0; i < 7; ++i) {
      double add = 0.0, kii = 0.0;
      for (int j = 0; j < 7; ++j) {
        double kij = 0.0;                      // K = P (P + R)^-1
        for (int k = 0; k < 7; ++k) kij += P[i][k] * Si[k][j];
        add += kij * res[j];
        if (j == i) kii = kij;
      }
      cur[i] += ok ? add : CUDART_NAN;       // np.linalg.inv raises on a singular matrix; here the state turns NaN
      pd[i] = (1.0 - kii) * P[i][i];           // element-wise (I - K o I) o P: diagonal survives, the rest becomes 0
    }
    {
      const double n = sqrt(cur[3] * cur[3] + cur[4] * cur[4] + cur[5] * cur[5] + cur[6] * cur[6]);
#pragma unroll
      for (int i = 3; i < 7; ++i) cur[i] /= n;
    }
#pragma unroll
    for (int i = 0; i < 7; ++i) filtered[o + i] = cur[i];
    if (p_diag)
#pragma unroll
      for (int i = 0; i < 7; ++i) p_diag[o + i] = pd[i];
    // ---- predict (filter.py:132-146, motion.py:27-53)
    double M[7] = {0, 0, 0, 0, 0, 0, 1};
    if (have_prev) {
      kf::relative_transform(prev, cur, M);
#pragma unroll
      for (int i = 0; i < 3; ++i) M[i] = (1 - a_t) * 0.0 + a_t * M[i];
      const double theta = acos(M[6]);         // dot((0,0,0,1), q); theta = 0 gives 0/0 = NaN like the reference
      const double s0 = sin((1 - a_r) * theta), s1 = sin(a_r * theta), sn = sin(theta);
      M[3] = (s0 * 0.0 + s1 * M[3]) / sn;
      M[4] = (s0 * 0.0 + s1 * M[4]) / sn;
      M[5] = (s0 * 0.0 + s1 * M[5]) / sn;
      M[6] = (s0 * 1.0 + s1 * M[6]) / sn;
    }
    double nxt[7];
    kf::apply_transform(cur, M, nxt);
    kf::jacobian(cur, M, nxt + 3, Sm);         // Sm reused as A
    {
      const double n = sqrt(nxt[3] * nxt[3] + nxt[4] * nxt[4] + nxt[5] * nxt[5] + nxt[6] * nxt[6]);
#pragma unroll
      for (int i = 3; i < 7; ++i) nxt[i] /= n;
    }
    for (int i = 0; i < 7; ++i)
      for (int j = 0; j < 7; ++j) {
        double v = i == j ? Q[i] : 0.0;        // A diag(pd) A^T + Q
        for (int k = 0; k < 7; ++k) v += Sm[i][k] * pd[k] * Sm[j][k];
        P[i][j] = v;
      }
    if (predicted)
#pragma unroll
      for (int i = 0; i < 7; ++i) predicted[o + i] = nxt[i];
#pragma unroll
    for (int i = 0; i < 7; ++i) { prev[i] = cur[i]; cur[i] = nxt[i]; }
    have_prev = true;
  }
}

}  // namespace plr2

extern "C" int plr2_kalman_filter_f64(const double *prior, const double *prior_var, const double *obs,
                                      const double *sensor_var, const double *motion_var, const int32_t *lengths,
                                      double a_t, double a_r, int sequences, int steps, double *filtered,
                                      double *p_diag, double *predicted, void *stream) {
  using namespace plr2;
  if (sequences < 0 || steps < 0) { set_error("plr2_kalman_filter_f64: bad shape S=%d T=%d", sequences, steps); return PLR2_E_SHAPE; }
  if (sequences == 0 || steps == 0) return 0;
  if (!prior || !prior_var || !obs || !sensor_var || !motion_var || !filtered) {
    set_error("plr2_kalman_filter_f64: null pointer");
    return PLR2_E_NULLPTR;
  }
  kalman_filter_kernel<<<(sequences + 63) / 64, 64, 0, (cudaStream_t)stream>>>(
      prior, prior_var, obs, sensor_var, motion_var, lengths, a_t, a_r, sequences, steps, filtered, p_diag, predicted);
  return check_launch("plr2_kalman_filter_f64");
}
