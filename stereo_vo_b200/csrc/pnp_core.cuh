// pnp_core.cuh - pose arithmetic shared by the device code of pnp_ransac.cu (fp64 throughout):
// Rodrigues exp / log, pinhole projection (no distortion: the reference passes zeros,
// src/image_processor.cpp:79), Gauss-Newton normal equations of the reprojection error under a
// left-multiplied rotation increment, a 6x6 Cholesky solve and a damped (Levenberg-Marquardt)
// step. Plain functions of their arguments, no memory traffic of their own.
//
// Pose convention is OpenCV's: x_cam = R * X_world + t, u = f * x / z + cx, v = f * y / z + cy.
#pragma once
#include <math.h>

#ifndef SVO_HD
#define SVO_HD __host__ __device__ __forceinline__
#endif

struct SvoPose {
    double R[9];   // row major
    double t[3];
};

// R = exp([w]x)  (cv::Rodrigues, vector -> matrix)
SVO_HD void svo_rodrigues_exp(const double w[3], double R[9]) {
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const double th = sqrt(th2);
    double a, b;                        // R = I + a [w]x + b [w]x^2
    if (th < 1e-8) {
        a = 1.0 - th2 / 6.0;
        b = 0.5 - th2 / 24.0;
    } else {
        a = sin(th) / th;
        b = (1.0 - cos(th)) / th2;
    }
    const double x = w[0], y = w[1], z = w[2];
    R[0] = 1.0 - b * (y * y + z * z); R[1] = -a * z + b * x * y;        R[2] = a * y + b * x * z;
    R[3] = a * z + b * x * y;         R[4] = 1.0 - b * (x * x + z * z); R[5] = -a * x + b * y * z;
    R[6] = -a * y + b * x * z;        R[7] = a * x + b * y * z;         R[8] = 1.0 - b * (x * x + y * y);
}

// w = log(R)  (cv::Rodrigues, matrix -> vector), R a rotation
SVO_HD void svo_rodrigues_log(const double R[9], double w[3]) {
    const double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    const double s = 0.5 * sqrt(rx * rx + ry * ry + rz * rz);      // sin(theta)
    double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);                   // cos(theta)
    c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
    const double th = atan2(s, c);
    if (s < 1e-10) {
        if (c > 0.0) {                  // theta ~ 0: w ~ vee(R - R^T) / 2
            w[0] = 0.5 * rx; w[1] = 0.5 * ry; w[2] = 0.5 * rz;
            return;
        }
        // theta ~ pi: axis from the diagonal of (R + I) / 2, signs from the off-diagonal sums
        double ax = sqrt(fmax((R[0] + 1.0) * 0.5, 0.0));
        double ay = sqrt(fmax((R[4] + 1.0) * 0.5, 0.0));
        double az = sqrt(fmax((R[8] + 1.0) * 0.5, 0.0));
        if (R[1] + R[3] < 0.0) ay = -ay;
        if (R[2] + R[6] < 0.0) az = -az;
        if (ax == 0.0 && R[5] + R[7] < 0.0) az = -az;
        const double n = sqrt(ax * ax + ay * ay + az * az);
        const double k = n > 0.0 ? th / n : 0.0;
        w[0] = ax * k; w[1] = ay * k; w[2] = az * k;
        return;
    }
    const double k = th / (2.0 * s);
    w[0] = rx * k; w[1] = ry * k; w[2] = rz * k;
}

// x_cam of a world point
SVO_HD void svo_to_camera(const SvoPose& P, const double X[3], double xc[3], double rx[3]) {
    rx[0] = P.R[0] * X[0] + P.R[1] * X[1] + P.R[2] * X[2];
    rx[1] = P.R[3] * X[0] + P.R[4] * X[1] + P.R[5] * X[2];
    rx[2] = P.R[6] * X[0] + P.R[7] * X[1] + P.R[8] * X[2];
    xc[0] = rx[0] + P.t[0]; xc[1] = rx[1] + P.t[1]; xc[2] = rx[2] + P.t[2];
}

// squared reprojection error of one point (1e30 behind the camera)
SVO_HD double svo_reproj_err2(const SvoPose& P, const double X[3], double u, double v, double f,
                              double cx, double cy) {
    double xc[3], rx[3];
    svo_to_camera(P, X, xc, rx);
    if (!(xc[2] > 1e-9)) return 1e30;
    const double iz = 1.0 / xc[2];
    const double du = f * xc[0] * iz + cx - u, dv = f * xc[1] * iz + cy - v;
    return du * du + dv * dv;
}

// One point's contribution to the normal equations: H (upper triangle, 21 entries row by row)
// += J^T J, g += J^T r, returns r^T r. Increment parametrisation: R <- exp([dw]x) R, t <- t + dt,
// unknowns (dw, dt). d x_cam / d dw = -[R X]x, d x_cam / d dt = I.
SVO_HD double svo_normal_eq_add(const SvoPose& P, const double X[3], double u, double v, double f,
                                double cx, double cy, double H[21], double g[6]) {
    double xc[3], rx[3];
    svo_to_camera(P, X, xc, rx);
    if (!(xc[2] > 1e-9)) return 1e30;
    const double iz = 1.0 / xc[2];
    const double px = xc[0] * iz, py = xc[1] * iz;
    const double ru = f * px + cx - u, rv = f * py + cy - v;
    // d(u, v) / d x_cam
    const double a0 = f * iz, a2 = -f * px * iz;       // du: (a0, 0, a2)
    const double b1 = f * iz, b2 = -f * py * iz;       // dv: (0, b1, b2)
    // columns: dw = rows of d x_cam/d dw = -[rx]x = [[0, rx2, -rx1], [-rx2, 0, rx0], [rx1, -rx0, 0]]
    double Ju[6], Jv[6];
    Ju[0] = a2 * rx[1];               Ju[1] = a0 * rx[2] - a2 * rx[0];  Ju[2] = -a0 * rx[1];
    Jv[0] = -b1 * rx[2] + b2 * rx[1]; Jv[1] = -b2 * rx[0];              Jv[2] = b1 * rx[0];
    Ju[3] = a0; Ju[4] = 0.0; Ju[5] = a2;
    Jv[3] = 0.0; Jv[4] = b1; Jv[5] = b2;
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = i; j < 6; ++j) H[k++] += Ju[i] * Ju[j] + Jv[i] * Jv[j];
        g[i] += Ju[i] * ru + Jv[i] * rv;
    }
    return ru * ru + rv * rv;
}

// solve (H + lambda diag(H)) d = -g for the 6 unknowns (H: upper triangle, 21 entries);
// false when the damped matrix is not positive definite
SVO_HD bool svo_solve6(const double H[21], const double g[6], double lambda, double d[6]) {
    double A[6][6];
    int k = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = i; j < 6; ++j) { A[i][j] = A[j][i] = H[k++]; }
#pragma unroll
    for (int i = 0; i < 6; ++i) A[i][i] += lambda * A[i][i] + 1e-12;
    // Cholesky A = L L^T in place (lower)
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double s = A[j][j];
#pragma unroll
        for (int p = 0; p < j; ++p) s -= A[j][p] * A[j][p];
        if (!(s > 0.0)) return false;
        const double l = sqrt(s);
        A[j][j] = l;
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            double t = A[i][j];
#pragma unroll
            for (int p = 0; p < j; ++p) t -= A[i][p] * A[j][p];
            A[i][j] = t / l;
        }
    }
    double y[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        double t = -g[i];
#pragma unroll
        for (int p = 0; p < i; ++p) t -= A[i][p] * y[p];
        y[i] = t / A[i][i];
    }
#pragma unroll
    for (int i = 5; i >= 0; --i) {
        double t = y[i];
#pragma unroll
        for (int p = i + 1; p < 6; ++p) t -= A[p][i] * d[p];
        d[i] = t / A[i][i];
    }
    return true;
}

// P' = (exp([d[0..2]]x) R, t + d[3..5])
SVO_HD void svo_pose_apply(const SvoPose& P, const double d[6], SvoPose& out) {
    double E[9];
    svo_rodrigues_exp(d, E);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            out.R[3 * i + j] = E[3 * i] * P.R[j] + E[3 * i + 1] * P.R[3 + j] + E[3 * i + 2] * P.R[6 + j];
    out.t[0] = P.t[0] + d[3]; out.t[1] = P.t[1] + d[4]; out.t[2] = P.t[2] + d[5];
}

// cv::RANSACUpdateNumIters(confidence, outlier ratio, model points, current bound)
SVO_HD int svo_ransac_update_iters(double p, double ep, int model_points, int max_iters) {
    p = fmax(p, 0.0); p = fmin(p, 1.0);
    ep = fmax(ep, 0.0); ep = fmin(ep, 1.0);
    double num = fmax(1.0 - p, 2.2250738585072014e-308);
    double denom = 1.0 - pow(1.0 - ep, (double)model_points);
    if (denom < 2.2250738585072014e-308) return 0;
    num = log(num);
    denom = log(denom);
    if (denom >= 0 || -num >= max_iters * (-denom)) return max_iters;
    return (int)rint(num / denom);
}
