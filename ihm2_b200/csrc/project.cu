// project.cu -- Cartesian pose -> Frenet state (s, n, psi) for a batch of cars.
//
// Reference: Track::project, src/ihm2/src/common/tracks.cpp:200-288 (window of table points around s_guess, closest
// point, neighbour chosen by the larger angle seen from the car, orthogonal projection on that segment, phi_ref from the
// linear interpolant anchored at the closest point, :191-198 / :169-170), wrap_to_pi of common/math.cpp:19-25, and the
// control node's lateral offset and heading error, src/ihm2/src/mpc_control_node.cpp:142-157.
//
// One thread per car: a binary search for the window (s_tol = 2 m is 4-8 table points), a scan for the closest point.
// The tables are shared by all cars of a track and stay in L2; per car 32 bytes come in and 24-52 go out.
#include "common.cuh"

namespace ihm2b {
namespace {
constexpr double PI_ = 3.14159265358979323846;

__device__ __forceinline__ double wrap_to_pi(double x) {
  double t = fmod(x + PI_, 2.0 * PI_);
  if (t < 0.0) t += 2.0 * PI_;
  return t - PI_;
}
// std::upper_bound(v, v + n, x) - v - 1
__device__ __forceinline__ int locate_index(const double* __restrict__ v, int n, double x) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (x < v[mid]) hi = mid; else lo = mid + 1;
  }
  return lo - 1;
}
__device__ __forceinline__ double angle3pt(double ax, double ay, double bx, double by, double cx, double cy) {
  return wrap_to_pi(atan2(cy - by, cx - bx) - atan2(ay - by, ax - bx));
}

__global__ void __launch_bounds__(128) project_kernel(long long B, int n_grid, const double* __restrict__ s_ref,
                                                      const double* __restrict__ X_ref, const double* __restrict__ Y_ref,
                                                      const double* __restrict__ phi_ref, const int32_t* __restrict__ track,
                                                      const double* __restrict__ pose, const double* __restrict__ s_guess,
                                                      double s_tol, double* __restrict__ frenet, double* __restrict__ ref,
                                                      int32_t* __restrict__ flag) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const size_t off = (size_t)(track ? track[i] : 0) * n_grid;
  const double *sr = s_ref + off, *Xr = X_ref + off, *Yr = Y_ref + off, *pr = phi_ref + off;
  const double cx = pose[i * 3], cy = pose[i * 3 + 1], cphi = pose[i * 3 + 2];
  const double sg = s_guess[i];
  const double s_low = fmax(sg - s_tol, sr[0]), s_up = fmin(sg + s_tol, sr[n_grid - 1]);
  int id_low = locate_index(sr, n_grid, s_low), id_up = locate_index(sr, n_grid, s_up);
  if (id_low > 0) --id_low;
  if (id_up < n_grid - 1) ++id_up;
  const int rows = id_up - id_low + 1;
  int id_min = 0;
  double best = __longlong_as_double(0x7ff0000000000000LL);
  for (int j = 0; j < rows; j++) {
    const double dx = Xr[id_low + j] - cx, dy = Yr[id_low + j] - cy;
    const double d = dx * dx + dy * dy;
    if (d < best) { best = d; id_min = j; }
  }
  const int id_prev = (id_min == 0) ? rows - 1 : id_min - 1;
  const int id_next = (id_min == rows - 1) ? 0 : id_min + 1;
  const double mx = Xr[id_low + id_min], my = Yr[id_low + id_min];
  const double px = Xr[id_low + id_prev], py = Yr[id_low + id_prev];
  const double nx = Xr[id_low + id_next], ny = Yr[id_low + id_next];
  const double angle_prev = fabs(angle3pt(mx, my, cx, cy, px, py)), angle_next = fabs(angle3pt(mx, my, cx, cy, nx, ny));
  const bool back = angle_prev > angle_next;
  const double ax = back ? px : mx, ay = back ? py : my, bx = back ? mx : nx, by = back ? my : ny;
  const double sa = sr[id_low + (back ? id_prev : id_min)], sb = sr[id_low + (back ? id_min : id_next)];
  const double dx = bx - ax, dy = by - ay;
  const double lambda = ((cx - ax) * dx + (cy - ay) * dy) / (dx * dx + dy * dy);
  const double s = sa + lambda * (sb - sa);
  const double Xp = ax + lambda * (bx - ax), Yp = ay + lambda * (by - ay);
  const int ind = min(id_min + id_low, n_grid - 2);   // the reference's coefficient table has n_grid - 1 rows
  const double phir = pr[ind] + (pr[ind + 1] - pr[ind]) / (sr[ind + 1] - sr[ind]) * (s - sr[ind]);
  const double rho = wrap_to_pi(phir);
  const double psi = wrap_to_pi(cphi - rho);
  const double e = hypot(Xp - cx, Yp - cy);
  double sn, cs;
  sincos(rho, &sn, &cs);
  const double tpr = (cy - Yp) * cs - (cx - Xp) * sn;
  frenet[i * 3] = s;
  frenet[i * 3 + 1] = e * (tpr > 0.0 ? 1.0 : -1.0);
  frenet[i * 3 + 2] = psi;
  if (ref) { ref[i * 3] = Xp; ref[i * 3 + 1] = Yp; ref[i * 3 + 2] = phir; }
  if (flag) flag[i] = fabs(tpr) < 1e-6;
}
}  // namespace

cudaError_t launch_project(long long B, int n_grid, const double* s_ref, const double* X_ref, const double* Y_ref,
                           const double* phi_ref, const int32_t* track, const double* pose, const double* s_guess,
                           double s_tol, double* frenet, double* ref, int32_t* flag, cudaStream_t stream) {
  project_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(B, n_grid, s_ref, X_ref, Y_ref, phi_ref, track, pose,
                                                                 s_guess, s_tol, frenet, ref, flag);
  return cudaGetLastError();
}
}  // namespace ihm2b
