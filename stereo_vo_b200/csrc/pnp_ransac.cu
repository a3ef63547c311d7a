// pnp_ransac: the reference's pose step, src/image_processor.cpp:75-92 -
//   cv::solvePnPRansac(world, image, K, zeros, rvec, tvec, useExtrinsicGuess = true, 100, 8.0,
//                      0.99, inliers);  cv::Rodrigues(rvec, R)
// - batched over sequences, one 128-thread block per sequence.
//
// What is kept from OpenCV's RANSAC (calib3d ptsetreg.cpp, solvepnp.cpp), because it decides
// WHICH points become inliers and how the final pose is defined:
//   * model size 5; a hypothesis wins only with strictly more inliers than the best so far (and
//     at least 5), hypotheses are consumed in index order and the loop stops at the bound
//     cv::RANSACUpdateNumIters(confidence, outlier ratio, 5, bound) recomputes after every new
//     best (so a better hypothesis beyond the bound is never looked at, as in OpenCV);
//   * inlier <=> squared reprojection error <= reproj_error^2 under the WINNING HYPOTHESIS (the
//     inlier list is not recomputed after refinement), indices ascending;
//   * the returned pose minimises the summed squared reprojection error over those inliers
//     (OpenCV: SOLVEPNP_ITERATIVE, Levenberg-Marquardt; here: LM on the same cost to
//     convergence, fp64), R = Rodrigues(rvec).
// What is different, and why results agree anyway: OpenCV draws its 5-point samples from
// cv::RNG(-1) and fits each with EPnP; here sample k of a call comes from a counter-based hash
// and is fitted by LM from the extrinsic guess (the reference always passes one). Any
// hypothesis that is correct to well under the 8-px threshold selects the same inlier set
// unless a point's error sits on the threshold, and the refined pose is the minimum of the same
// cost over the same points: BASELINE.md states the tolerance (identical inlier sets on the
// synthetic sequences, |dt| <= 1 mm, rotation <= 1e-4 rad).
// All hypotheses of a call are fitted and scored in parallel (thread per fit, warp per score);
// OpenCV's sequential early stop is then replayed over the counts by one thread.
#include "pnp_core.cuh"
#include "svo_common.cuh"

namespace {

constexpr int PNP_THREADS = 128;
constexpr int PNP_MAX_HYP = 128;          // hypotheses fitted per call (iterations <= this)
constexpr int PNP_MODEL = 5;
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ unsigned pnp_hash(unsigned a, unsigned b, unsigned c) {
    // 3-word mix (lowbias32 rounds): the sample of (call salt, hypothesis, draw)
    unsigned x = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u ^ (c + 0x165667B1u) * 0xC2B2AE3Du;
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

struct PnpShared {
    SvoPose hyp[PNP_MAX_HYP];
    int count[PNP_MAX_HYP];
    double red[PNP_THREADS / 32][28];
    SvoPose cur, trial;
    double step[6];
    double cost, trial_cost, lambda;
    int best, n_inl, flag, tried;
};

// block-wide sum of 28 doubles (H, g, cost): result in sh.red[0][*] after the second barrier
__device__ __forceinline__ void block_sum28(double v[28], PnpShared& sh, int tid) {
#pragma unroll
    for (int k = 0; k < 28; ++k) {
        double x = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(FULL, x, o);
        v[k] = x;
    }
    if ((tid & 31) == 0)
        for (int k = 0; k < 28; ++k) sh.red[tid >> 5][k] = v[k];
    __syncthreads();
    if (tid < 28) {
        double s = 0.0;
        for (int w = 0; w < PNP_THREADS / 32; ++w) s += sh.red[w][tid];
        sh.red[0][tid] = s;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(PNP_THREADS)
pnp_ransac_kernel(const float* __restrict__ world, const float2* __restrict__ image, int stride,
                  const int* __restrict__ counts, const double* __restrict__ guess, double f,
                  double cx, double cy, int iterations, double thr2, double confidence,
                  unsigned salt, int staged, SvoPnpResult* __restrict__ results,
                  int* __restrict__ inliers) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double s_part[PNP_THREADS / 32];
    PnpShared& sh = *reinterpret_cast<PnpShared*>(smem_raw);
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const int n = min(counts[b], stride);
    SvoPnpResult* res = results + b;
    int* inl = inliers + (size_t)b * stride;
    // the points of this sequence: staged in shared memory when they fit (the reference's 400
    // features: 8 KB), else read in place
    const float* s_world = world + (size_t)b * stride * 3;                        // [n][3]
    const float2* s_img = image + (size_t)b * stride;                             // [n]
    if (staged) {
        float* sw = reinterpret_cast<float*>(smem_raw + sizeof(PnpShared));
        float2* si = reinterpret_cast<float2*>(sw + 3 * (size_t)stride);
        for (int i = tid; i < n; i += PNP_THREADS) {
            sw[3 * i + 0] = s_world[3 * i + 0];
            sw[3 * i + 1] = s_world[3 * i + 1];
            sw[3 * i + 2] = s_world[3 * i + 2];
            si[i] = s_img[i];
        }
        s_world = sw;
        s_img = si;
    }
    SvoPose P0;
    {
        const double* gq = guess + (size_t)b * 6;
        const double w[3] = {gq[0], gq[1], gq[2]};
        svo_rodrigues_exp(w, P0.R);
        P0.t[0] = gq[3]; P0.t[1] = gq[4]; P0.t[2] = gq[5];
    }
    __syncthreads();
    const int H = min(iterations, PNP_MAX_HYP);
    if (n < PNP_MODEL || H <= 0) {            // OpenCV: fewer points than the model -> failure
        if (tid == 0) {
            res->ok = 0; res->n_inliers = 0; res->best_hypothesis = -1; res->hypotheses_tried = 0;
            for (int k = 0; k < 3; ++k) { res->rvec[k] = guess[(size_t)b * 6 + k]; res->tvec[k] = P0.t[k]; }
            for (int k = 0; k < 9; ++k) res->rmat[k] = P0.R[k];
        }
        return;
    }

    // ---- phase A: thread h fits hypothesis h to its 5-point sample (LM from the guess)
    if (tid < H) {
        int idx[PNP_MODEL];
        for (int j = 0; j < PNP_MODEL; ++j) {
            unsigned draw = 0;
            for (;;) {
                const int c = (int)(pnp_hash(salt + (unsigned)b * 0x632BE5ABu, (unsigned)tid,
                                             (unsigned)(j + PNP_MODEL * draw)) % (unsigned)n);
                bool dup = false;
                for (int q = 0; q < j; ++q) dup |= idx[q] == c;
                if (!dup) { idx[j] = c; break; }
                ++draw;
            }
        }
        SvoPose P = P0;
        double lambda = 1e-3, cost = 1e300;
        for (int it = 0; it < 12; ++it) {
            double Hm[21], g[6], c = 0.0;
            for (int k = 0; k < 21; ++k) Hm[k] = 0.0;
            for (int k = 0; k < 6; ++k) g[k] = 0.0;
            for (int j = 0; j < PNP_MODEL; ++j) {
                const double X[3] = {s_world[3 * idx[j]], s_world[3 * idx[j] + 1], s_world[3 * idx[j] + 2]};
                c += svo_normal_eq_add(P, X, s_img[idx[j]].x, s_img[idx[j]].y, f, cx, cy, Hm, g);
            }
            if (it == 0) cost = c;
            bool moved = false;
            for (int tries = 0; tries < 6 && !moved; ++tries) {
                double d[6];
                if (svo_solve6(Hm, g, lambda, d)) {
                    SvoPose T;
                    svo_pose_apply(P, d, T);
                    double tc = 0.0;
                    for (int j = 0; j < PNP_MODEL; ++j) {
                        const double X[3] = {s_world[3 * idx[j]], s_world[3 * idx[j] + 1], s_world[3 * idx[j] + 2]};
                        tc += svo_reproj_err2(T, X, s_img[idx[j]].x, s_img[idx[j]].y, f, cx, cy);
                    }
                    if (tc <= cost) {
                        const bool small = cost - tc <= 1e-12 * (1.0 + cost);
                        P = T; cost = tc; lambda = fmax(lambda * 0.1, 1e-9); moved = true;
                        if (small) it = 1000;
                        break;
                    }
                }
                lambda *= 10.0;
            }
            if (!moved) break;
        }
        sh.hyp[tid] = P;
    }
    __syncthreads();

    // ---- phase B: inlier count of every hypothesis (a warp per hypothesis, lanes over points)
    for (int h = wrp; h < H; h += PNP_THREADS / 32) {
        const SvoPose P = sh.hyp[h];
        int c = 0;
        for (int i = lane; i < n; i += 32) {
            const double X[3] = {s_world[3 * i], s_world[3 * i + 1], s_world[3 * i + 2]};
            c += svo_reproj_err2(P, X, s_img[i].x, s_img[i].y, f, cx, cy) <= thr2;
        }
        c = __reduce_add_sync(FULL, c);
        if (lane == 0) sh.count[h] = c;
    }
    __syncthreads();

    // ---- OpenCV's sequential loop replayed over the counts (RANSACPointSetRegistrator::run)
    if (tid == 0) {
        int niters = H, best = -1, best_count = 0, h = 0;
        for (; h < niters; ++h) {
            const int c = sh.count[h];
            if (c > max(best_count, PNP_MODEL - 1)) {
                best = h;
                best_count = c;
                niters = svo_ransac_update_iters(confidence, (double)(n - c) / n, PNP_MODEL, niters);
            }
        }
        sh.best = best;
        sh.tried = h;
        sh.n_inl = 0;
    }
    __syncthreads();
    const int best = sh.best;
    if (best < 0) {
        if (tid == 0) {
            res->ok = 0; res->n_inliers = 0; res->best_hypothesis = -1; res->hypotheses_tried = sh.tried;
            for (int k = 0; k < 3; ++k) { res->rvec[k] = guess[(size_t)b * 6 + k]; res->tvec[k] = P0.t[k]; }
            for (int k = 0; k < 9; ++k) res->rmat[k] = P0.R[k];
        }
        return;
    }

    // ---- inliers of the winning hypothesis, ascending (ordered compaction, warp 0)
    if (wrp == 0) {
        const SvoPose P = sh.hyp[best];
        int base = 0;
        for (int i0 = 0; i0 < n; i0 += 32) {
            const int i = i0 + lane;
            bool in = false;
            if (i < n) {
                const double X[3] = {s_world[3 * i], s_world[3 * i + 1], s_world[3 * i + 2]};
                in = svo_reproj_err2(P, X, s_img[i].x, s_img[i].y, f, cx, cy) <= thr2;
            }
            const unsigned m = __ballot_sync(FULL, in);
            if (in) inl[base + __popc(m & ((1u << lane) - 1u))] = i;
            base += __popc(m);
        }
        if (lane == 0) {
            sh.n_inl = base;
            sh.cur = P;
            sh.lambda = 1e-3;
        }
    }
    __syncthreads();
    const int n_inl = sh.n_inl;

    // ---- refinement: LM over the inliers, every thread a slice of the points
    auto gather = [&](const SvoPose& P, bool with_eq, double acc[28]) {
        for (int k = 0; k < 28; ++k) acc[k] = 0.0;
        for (int q = tid; q < n_inl; q += PNP_THREADS) {
            const int i = inl[q];
            const double X[3] = {s_world[3 * i], s_world[3 * i + 1], s_world[3 * i + 2]};
            if (with_eq) acc[27] += svo_normal_eq_add(P, X, s_img[i].x, s_img[i].y, f, cx, cy, acc, acc + 21);
            else acc[27] += svo_reproj_err2(P, X, s_img[i].x, s_img[i].y, f, cx, cy);
        }
    };
    __threadfence_block();
    for (int it = 0; it < 20; ++it) {
        double acc[28];
        const SvoPose P = sh.cur;
        gather(P, true, acc);
        block_sum28(acc, sh, tid);
        if (tid == 0) {
            sh.cost = sh.red[0][27];
            sh.flag = 0;
        }
        __syncthreads();
        bool moved = false;
        for (int tries = 0; tries < 8; ++tries) {
            if (tid == 0) {
                double d[6];
                const bool ok = svo_solve6(sh.red[0], sh.red[0] + 21, sh.lambda, d);
                if (ok) svo_pose_apply(P, d, sh.trial);
                double dn = 0.0;
                for (int k = 0; k < 6; ++k) { sh.step[k] = d[k]; dn += ok ? d[k] * d[k] : 0.0; }
                sh.flag = ok ? (dn < 1e-28 ? 2 : 1) : 0;
            }
            __syncthreads();
            const int flag = sh.flag;
            if (flag == 2) break;                         // step below resolution: converged
            if (flag == 1) {
                double tacc[28];
                const SvoPose T = sh.trial;
                gather(T, false, tacc);
                // only the cost entry matters; reduce it alone
                double x = tacc[27];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(FULL, x, o);
                if (lane == 0) s_part[wrp] = x;
                __syncthreads();
                if (tid == 0) {
                    double tc = 0.0;
                    for (int w = 0; w < PNP_THREADS / 32; ++w) tc += s_part[w];
                    sh.trial_cost = tc;
                    if (tc <= sh.cost) {
                        sh.cur = sh.trial;
                        sh.lambda = fmax(sh.lambda * 0.1, 1e-12);
                        sh.flag = (sh.cost - tc <= 1e-14 * (1.0 + sh.cost)) ? 4 : 3;   // 4: converged
                        sh.cost = tc;
                    } else {
                        sh.lambda *= 10.0;
                        sh.flag = 5;
                    }
                }
                __syncthreads();
                if (sh.flag == 3 || sh.flag == 4) { moved = true; break; }
            } else {
                if (tid == 0) sh.lambda *= 10.0;
                __syncthreads();
            }
        }
        const int flag = sh.flag;
        __syncthreads();
        if (!moved || flag == 4) break;
    }

    if (tid == 0) {
        const SvoPose P = sh.cur;
        double w[3];
        svo_rodrigues_log(P.R, w);
        for (int k = 0; k < 3; ++k) { res->rvec[k] = w[k]; res->tvec[k] = P.t[k]; }
        for (int k = 0; k < 9; ++k) res->rmat[k] = P.R[k];
        res->n_inliers = n_inl;
        res->ok = 1;
        res->best_hypothesis = best;
        res->hypotheses_tried = sh.tried;
    }
}

}  // namespace

static size_t pnp_smem_bytes(int stride, bool staged) {
    return sizeof(PnpShared) + (staged ? (size_t)stride * (3 * sizeof(float) + sizeof(float2)) : 0) + 16;
}

cudaError_t svo_launch_pnp_ransac(const float* world, const float2* image, int stride,
                                  const int* counts, const double* guess, float focal, float cx,
                                  float cy, int iterations, float reproj_error, double confidence,
                                  unsigned salt, SvoPnpResult* results, int* inliers, int n_call,
                                  cudaStream_t st) {
    if (n_call <= 0) return cudaSuccess;
    const bool staged = pnp_smem_bytes(stride, true) <= 48 * 1024;
    const size_t smem = pnp_smem_bytes(stride, staged);
    pnp_ransac_kernel<<<n_call, PNP_THREADS, smem, st>>>(
        world, image, stride, counts, guess, (double)focal, (double)cx, (double)cy, iterations,
        (double)reproj_error * (double)reproj_error, confidence, salt, staged ? 1 : 0, results,
        inliers);
    return cudaGetLastError();
}
