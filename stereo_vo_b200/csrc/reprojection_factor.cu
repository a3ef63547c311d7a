// reprojection_factor: ReprojectionFactor::Evaluate of the reference
// (src/reprojection_factor.cpp:10-88, ceres::SizedCostFunction<2, 7, 3>) for every observation
// of a bundle-adjustment window in one launch: <= 400 observations x 5
// keyframes (src/vo_node.cpp:36, src/bundle_adjuster.hpp:75), one thread per observation, fp64.
//
//   pose  = (qw, qx, qy, qz, tx, ty, tz): world frame with respect to the camera (:16-17)
//   gamma = LR p / |q|^2 + t,  LR = v v^T + ([v]x + qw I)^2 = |q|^2 R(q)            (:26-35)
//   r     = (f gamma_x / gamma_z + cx - ox,  f gamma_y / gamma_z + cy - oy)        (:37-40)
//   d r / d pose (2 x 7 row major) and d r / d point (2 x 3): the reference's MATLAB-generated
//   expressions (:62-85) are the derivatives of exactly this function, normalisation included;
//   here they come from the chain rule through gamma (LR p = (qw^2 - |v|^2) p + 2 v (v.p) +
//   2 qw (v x p)). Pinned against the reference's expressions by tests/golden/
//   reprojection_factor.npz (<= 1e-12 relative).
// Latency-bound by construction (a few hundred observations); the launch exists so that a GPU
// Schur solver - or the host Ceres problem - can take residuals and Jacobians of a whole window
// without a per-factor virtual call.
#include "svo_common.cuh"

namespace {

__global__ void __launch_bounds__(128)
reprojection_factor_kernel(const double* __restrict__ poses, const double* __restrict__ points,
                           const int* __restrict__ pose_idx, const int* __restrict__ point_idx,
                           const double* __restrict__ obs, int n_obs, double f, double cx, double cy,
                           double* __restrict__ resid, double* __restrict__ jac_pose,
                           double* __restrict__ jac_point) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_obs) return;
    const double* q = poses + (size_t)pose_idx[i] * 7;
    const double* pp = points + (size_t)point_idx[i] * 3;
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    const double p0 = pp[0], p1 = pp[1], p2 = pp[2];
    const double n = w * w + x * x + y * y + z * z, in = 1.0 / n;
    const double vv = x * x + y * y + z * z, vp = x * p0 + y * p1 + z * p2;
    // v x p
    const double c0 = y * p2 - z * p1, c1 = z * p0 - x * p2, c2 = x * p1 - y * p0;
    const double s = w * w - vv;
    const double m0 = s * p0 + 2.0 * x * vp + 2.0 * w * c0;      // LR p
    const double m1 = s * p1 + 2.0 * y * vp + 2.0 * w * c1;
    const double m2 = s * p2 + 2.0 * z * vp + 2.0 * w * c2;
    const double g0 = m0 * in + q[4], g1 = m1 * in + q[5], g2 = m2 * in + q[6];
    const double psi = 1.0 / g2;
    if (resid) {
        resid[2 * i + 0] = f * g0 * psi + cx - obs[2 * i + 0];
        resid[2 * i + 1] = f * g1 * psi + cy - obs[2 * i + 1];
    }
    // d r / d gamma = [a 0 b0; 0 a b1]
    const double a = f * psi, b0 = -f * g0 * psi * psi, b1 = -f * g1 * psi * psi;
    if (jac_pose) {
        // columns of d(LR p)/dq: qw, then qx, qy, qz (e_i x p = rows of -[p]x)
        double d[3][4];
        d[0][0] = 2.0 * (w * p0 + c0); d[1][0] = 2.0 * (w * p1 + c1); d[2][0] = 2.0 * (w * p2 + c2);
        // d/dx: -2 x p + 2 e_x vp + 2 v p0 + 2 w (e_x x p),  e_x x p = (0, -p2, p1)
        d[0][1] = -2.0 * x * p0 + 2.0 * vp + 2.0 * x * p0;          d[1][1] = -2.0 * x * p1 + 2.0 * y * p0 - 2.0 * w * p2; d[2][1] = -2.0 * x * p2 + 2.0 * z * p0 + 2.0 * w * p1;
        // e_y x p = (p2, 0, -p0)
        d[0][2] = -2.0 * y * p0 + 2.0 * x * p1 + 2.0 * w * p2;      d[1][2] = -2.0 * y * p1 + 2.0 * vp + 2.0 * y * p1;     d[2][2] = -2.0 * y * p2 + 2.0 * z * p1 - 2.0 * w * p0;
        // e_z x p = (-p1, p0, 0)
        d[0][3] = -2.0 * z * p0 + 2.0 * x * p2 - 2.0 * w * p1;      d[1][3] = -2.0 * z * p1 + 2.0 * y * p2 + 2.0 * w * p0; d[2][3] = -2.0 * z * p2 + 2.0 * vp + 2.0 * z * p2;
        const double qq[4] = {w, x, y, z};
        double* jp = jac_pose + (size_t)i * 14;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const double k2 = 2.0 * qq[k] * in * in;          // the 1/|q|^2 factor differentiated
            const double e0 = d[0][k] * in - m0 * k2, e1 = d[1][k] * in - m1 * k2,
                         e2 = d[2][k] * in - m2 * k2;
            jp[k] = a * e0 + b0 * e2;
            jp[7 + k] = a * e1 + b1 * e2;
        }
        jp[4] = a;   jp[5] = 0.0; jp[6] = b0;
        jp[11] = 0.0; jp[12] = a;  jp[13] = b1;
    }
    if (jac_point) {
        // R = LR / |q|^2
        const double r00 = (s + 2.0 * x * x) * in, r01 = 2.0 * (x * y - w * z) * in, r02 = 2.0 * (x * z + w * y) * in;
        const double r10 = 2.0 * (x * y + w * z) * in, r11 = (s + 2.0 * y * y) * in, r12 = 2.0 * (y * z - w * x) * in;
        const double r20 = 2.0 * (x * z - w * y) * in, r21 = 2.0 * (y * z + w * x) * in, r22 = (s + 2.0 * z * z) * in;
        double* jx = jac_point + (size_t)i * 6;
        jx[0] = a * r00 + b0 * r20; jx[1] = a * r01 + b0 * r21; jx[2] = a * r02 + b0 * r22;
        jx[3] = a * r10 + b1 * r20; jx[4] = a * r11 + b1 * r21; jx[5] = a * r12 + b1 * r22;
    }
}

}  // namespace

cudaError_t svo_launch_reprojection_factor(const double* poses, const double* points,
                                           const int* pose_idx, const int* point_idx,
                                           const double* obs, int n_obs, double f, double cx,
                                           double cy, double* resid, double* jac_pose,
                                           double* jac_point, cudaStream_t st) {
    if (n_obs <= 0) return cudaSuccess;
    reprojection_factor_kernel<<<(n_obs + 127) / 128, 128, 0, st>>>(
        poses, points, pose_idx, point_idx, obs, n_obs, f, cx, cy, resid, jac_pose, jac_point);
    return cudaGetLastError();
}
