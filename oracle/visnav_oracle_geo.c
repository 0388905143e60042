/* visnav_oracle_geo.c -- CPU restatement of the geometry helpers that follow matching.
 * TEST INFRASTRUCTURE ONLY: linked into oracle/liboracle.so; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline leg may call it.  Parity status: PINNED against the reference's own code
 * (oracle/_ref/libvisnav_ref_geo.so, built from camera_models.h / matching_utils.h / vo_utils.h where
 * they lie) by tests/test_oracle_golden.py and the committed fixture tests/golden/geometry.npz.
 *
 * Follows, line by line:
 *   camera models  include/visnav/camera_models.h  pinhole :75-117, eucm :158-212, ds :246-305, kb4 :341-439
 *   findInliersEssential  include/visnav/matching_utils.h:63-88
 *   project_landmarks     include/visnav/vo_utils.h:48-81 (pose inverse and point action:
 *                         thirdparty/Sophus/sophus/se3.hpp:208-211, so3.hpp:229-231, :362-371)
 * Floating point: doubles throughout, expressions in the reference's order; compiled with
 * -ffp-contract=off whereas the reference's -O3 build may fuse multiply-adds, so agreement with the
 * reference is to a few ulps (tests state the tolerance), not bit-exact.
 */
#include <math.h>
#include <stdint.h>

enum { CAM_PINHOLE = 0, CAM_EUCM = 1, CAM_DS = 2, CAM_KB4 = 3 };

static void cam_project(int model, const double* P, const double* p, double* uv) {
  const double fx = P[0], fy = P[1], cx = P[2], cy = P[3];
  const double x = p[0], y = p[1], z = p[2];
  switch (model) {
    case CAM_PINHOLE: /* camera_models.h:89-90 */
      uv[0] = fx * x / z + cx;
      uv[1] = fy * y / z + cy;
      break;
    case CAM_EUCM: { /* :173-175 */
      const double alpha = P[4], beta = P[5];
      const double d = sqrt(beta * (x * x + y * y) + z * z);
      uv[0] = fx * x / (alpha * d + (1.0 - alpha) * z) + cx;
      uv[1] = fy * y / (alpha * d + (1.0 - alpha) * z) + cy;
      break;
    }
    case CAM_DS: { /* :261-267 */
      const double xi = P[4], alpha = P[5];
      const double d1 = sqrt(x * x + y * y + z * z);
      const double d2 = sqrt(x * x + y * y + (xi * d1 + z) * (xi * d1 + z));
      uv[0] = fx * x / (alpha * d2 + (1.0 - alpha) * (xi * d1 + z)) + cx;
      uv[1] = fy * y / (alpha * d2 + (1.0 - alpha) * (xi * d1 + z)) + cy;
      break;
    }
    default: { /* kb4 :358-372 */
      const double k1 = P[4], k2 = P[5], k3 = P[6], k4 = P[7];
      const double r = sqrt(x * x + y * y);
      const double t = atan2(r, z);
      const double d = t + k1 * t * t * t + k2 * t * t * t * t * t + k3 * t * t * t * t * t * t * t +
                       k4 * t * t * t * t * t * t * t * t * t;
      if (r == 0.0) {
        uv[0] = cx;
        uv[1] = cy;
      } else {
        uv[0] = fx * d * x / r + cx;
        uv[1] = fy * d * y / r + cy;
      }
    }
  }
}

static void cam_unproject(int model, const double* P, const double* uv, double* out) {
  const double fx = P[0], fy = P[1], cx = P[2], cy = P[3];
  const double mx = (uv[0] - cx) / fx, my = (uv[1] - cy) / fy;
  switch (model) {
    case CAM_PINHOLE: { /* :106-115 */
      const double s = 1.0 / sqrt(mx * mx + my * my + 1.0);
      out[0] = mx * s;
      out[1] = my * s;
      out[2] = 1.0 * s;
      break;
    }
    case CAM_EUCM: { /* :195-207 */
      const double alpha = P[4], beta = P[5];
      const double rr = mx * mx + my * my;
      const double mz = (1.0 - beta * alpha * alpha * rr) /
                        (alpha * sqrt(1.0 - (2.0 * alpha - 1.0) * beta * rr) + (1.0 - alpha));
      const double s = 1.0 / sqrt(mx * mx + my * my + mz * mz);
      out[0] = mx * s;
      out[1] = my * s;
      out[2] = mz * s;
      break;
    }
    case CAM_DS: { /* :287-302 */
      const double xi = P[4], alpha = P[5];
      const double rr = mx * mx + my * my;
      const double mz = (1.0 - alpha * alpha * rr) / (alpha * sqrt(1.0 - (2.0 * alpha - 1.0) * rr) + 1.0 - alpha);
      const double s = (mz * xi + sqrt(mz * mz + (1.0 - xi * xi) * rr)) / (mz * mz + rr);
      out[0] = mx * s;
      out[1] = my * s;
      out[2] = mz * s - xi;
      break;
    }
    default: { /* kb4 :391-436: five Newton steps from theta = 0 */
      const double k1 = P[4], k2 = P[5], k3 = P[6], k4 = P[7];
      const double ru = sqrt(mx * mx + my * my);
      double x = 0.0;
      for (int itr = 0; itr < 5; itr++) {
        const double d = x + k1 * x * x * x + k2 * x * x * x * x * x + k3 * x * x * x * x * x * x * x +
                         k4 * x * x * x * x * x * x * x * x * x - ru;
        const double diff = 1.0 + 3.0 * k1 * x * x + 5.0 * k2 * x * x * x * x + 7.0 * k3 * x * x * x * x * x * x +
                            9.0 * k4 * x * x * x * x * x * x * x * x;
        x = x - d / diff;
      }
      if (ru == 0.0) {
        out[0] = 0.0;
        out[1] = 0.0;
        out[2] = cos(x);
      } else {
        out[0] = sin(x) * mx / ru;
        out[1] = sin(x) * my / ru;
        out[2] = cos(x);
      }
    }
  }
}

void vso_cam_project(int model, const double* param, const double* xyz, int n, double* uv) {
  for (int i = 0; i < n; i++) cam_project(model, param, xyz + 3 * i, uv + 2 * i);
}

void vso_cam_unproject(int model, const double* param, const double* uv, int n, double* xyz) {
  for (int i = 0; i < n; i++) cam_unproject(model, param, uv + 2 * i, xyz + 3 * i);
}

/* matching_utils.h:63-88.  E row-major.  err_out (nullable): p0^T E p1 per match.  Returns #inliers. */
int vso_find_inliers_essential(const double* xy1, int n1, const double* xy2, int n2, int model1,
                               const double* param1, int model2, const double* param2, const double* E, double thr,
                               const int32_t* matches, int n_matches, int32_t* inliers, double* err_out) {
  int m = 0;
  (void)n1;
  (void)n2;
  for (int j = 0; j < n_matches; j++) {
    double a[3], b[3];
    cam_unproject(model1, param1, xy1 + 2 * matches[2 * j], a);
    cam_unproject(model2, param2, xy2 + 2 * matches[2 * j + 1], b);
    /* (p0^T E) p1: Eigen evaluates the row vector first, then the dot product */
    const double r0 = a[0] * E[0] + a[1] * E[3] + a[2] * E[6];
    const double r1 = a[0] * E[1] + a[1] * E[4] + a[2] * E[7];
    const double r2 = a[0] * E[2] + a[1] * E[5] + a[2] * E[8];
    const double err = r0 * b[0] + r1 * b[1] + r2 * b[2];
    if (err_out) err_out[j] = err;
    if (!(fabs(err) > thr)) { /* :80-84 */
      inliers[2 * m] = matches[2 * j];
      inliers[2 * m + 1] = matches[2 * j + 1];
      m++;
    }
  }
  return m;
}

/* so3.hpp:362-371: p + w*uv + qv x uv with uv = 2 (qv x p) */
static void quat_rotate(const double* q, const double* p, double* out) {
  double uv0 = q[1] * p[2] - q[2] * p[1], uv1 = q[2] * p[0] - q[0] * p[2], uv2 = q[0] * p[1] - q[1] * p[0];
  uv0 += uv0;
  uv1 += uv1;
  uv2 += uv2;
  out[0] = p[0] + q[3] * uv0 + (q[1] * uv2 - q[2] * uv1);
  out[1] = p[1] + q[3] * uv1 + (q[2] * uv0 - q[0] * uv2);
  out[2] = p[2] + q[3] * uv2 + (q[0] * uv1 - q[1] * uv0);
}

/* vo_utils.h:48-81.  pose = Sophus::SE3d::data(): qx qy qz qw tx ty tz (T_w_c).  Landmark i is reported by
 * its index; output order = ascending index.  Returns the number of projected points. */
int vso_project_landmarks(const double* pose, int model, const double* param, int width, int height,
                          const double* xyz, int n, double cam_z_threshold, double* proj_xy, int32_t* proj_idx) {
  /* se3.hpp:208-211: inverse = (conj(q) renormalised, conj(q) * (-t)) */
  const double len = sqrt(pose[0] * pose[0] + pose[1] * pose[1] + pose[2] * pose[2] + pose[3] * pose[3]);
  const double qi[4] = {-pose[0] / len, -pose[1] / len, -pose[2] / len, pose[3] / len};
  const double nt[3] = {pose[4] * -1.0, pose[5] * -1.0, pose[6] * -1.0};
  double ti[3];
  quat_rotate(qi, nt, ti);
  int m = 0;
  for (int i = 0; i < n; i++) {
    double pc[3], uv[2];
    quat_rotate(qi, xyz + 3 * i, pc);
    pc[0] += ti[0];
    pc[1] += ti[1];
    pc[2] += ti[2];
    if (pc[2] < cam_z_threshold) continue; /* :64 */
    cam_project(model, param, pc, uv);
    if (uv[0] > width || uv[1] > height || uv[0] < 0 || uv[1] < 0) continue; /* :69-70 */
    proj_xy[2 * m] = uv[0];
    proj_xy[2 * m + 1] = uv[1];
    proj_idx[m] = i;
    m++;
  }
  return m;
}
