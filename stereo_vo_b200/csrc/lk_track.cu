// lk_track: cv::calcOpticalFlowPyrLK(prev, next, pts, Size(21,21), maxLevel 3,
// {COUNT+EPS, 30, 0.01}, flags 0, minEigThreshold 1e-2) - reference call sites
// src/feature_tracker.cpp:23-26 (forward) and :32-35 (backward); arithmetic pinned in
// oracle/restate.py:lk_track.
//
// One block of NW warps per key-point, walking pyramid levels top -> 0 like LKTrackerInvoker:
//   * the 24x24 raw patch of I around the window is staged in shared memory (reflect-101, the
//     21-px border OpenCV keeps around each level), unnormalised Scharr derivatives are built
//     on the fly for the 22x22 integer grid (zero outside the image, as OpenCV's constant
//     border on the derivative image), then each thread interpolates its share of the 441
//     window pixels with the 14-bit fixed-point bilinear weights (I*32, Ix, Iy as int16 range);
//   * A11/A12/A22 and, per iteration, b1/b2 are EXACT integer window sums (per-thread int32,
//     warp REDUX on 16-bit halves, int64 across warps) converted once to f32 - cv2 accumulates
//     in f32 SIMD lanes, the two agree to <= 1e-4 px (tolerance 0.01 px, BASELINE.json);
//   * a 28x28 region of J is staged per level and re-staged only when the window leaves it, so
//     an iteration is shared-memory + integer work only; all threads carry the scalar
//     Gauss-Newton state redundantly, so control flow is uniform and needs no broadcast.
// Issue/shared-memory bound (<= 400 points x 4 levels x <= 30 iterations), not HBM-bound.
#include "svo_common.cuh"

namespace {

constexpr int WIN = SVO_LK_WIN;            // 21
constexpr int WIN2 = WIN * WIN;            // 441
constexpr int IP = WIN + 3;                // raw I patch edge (24)
constexpr int DG = WIN + 1;                // derivative grid edge (22)
constexpr int JM = 3;                      // margin of the staged J region
constexpr int JR = WIN + 1 + 2 * JM;       // staged J region edge (28)
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ long long warp_sum_exact(int v) {
    int lo = __reduce_add_sync(FULL, v & 0xffff);
    int hi = __reduce_add_sync(FULL, v >> 16);
    return ((long long)hi << 16) + (long long)lo;
}

__device__ __forceinline__ int clamp_reflect(int i, int n) {
    i = svo_reflect101(i, n);
    return min(max(i, 0), n - 1);
}

__device__ __forceinline__ void lk_weights(float a, float b, int& w00, int& w01, int& w10,
                                           int& w11) {
    const float s = 16384.f;                // 1 << W_BITS
    float oma = __fsub_rn(1.f, a), omb = __fsub_rn(1.f, b);
    w00 = __float2int_rn(__fmul_rn(__fmul_rn(oma, omb), s));
    w01 = __float2int_rn(__fmul_rn(__fmul_rn(a, omb), s));
    w10 = __float2int_rn(__fmul_rn(__fmul_rn(oma, b), s));
    w11 = 16384 - w00 - w01 - w10;
}

template <int NW>
__global__ void __launch_bounds__(NW * 32)
lk_track_kernel(const uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride,
                const float2* __restrict__ in_pts, size_t in_stride, int in_buffered,
                float2* __restrict__ out_pts, unsigned char* __restrict__ status, int out_stride,
                int direction, int max_iters, double eps2, double min_eig_thr, int max_pts,
                const int* __restrict__ gate, int gate_min,
                const SvoSeqCall* __restrict__ call) {
    constexpr int NT = NW * 32;
    constexpr int NPT = (WIN2 + NT - 1) / NT;

    __shared__ unsigned char s_I[IP][IP];
    __shared__ short s_dx[DG][DG];
    __shared__ short s_dy[DG][DG];
    __shared__ unsigned char s_J[JR][JR];
    __shared__ long long s_red[2][NW][3];

    const SvoSeqCall cd = call[blockIdx.y];
    const int pi = blockIdx.x;
    if (pi >= cd.n_feat || (cd.flags & 1) == 0) return;
    if (gate && gate[cd.seq] < gate_min) return;   // < 4 corners: the reference returns first
    const int tid = threadIdx.x;
    const int wrp = tid >> 5;

    const uint8_t* seq_base = pyr + (size_t)cd.seq * seq_stride;
    const int slot_i = direction == 0 ? (1 - cd.cur_slot) : cd.cur_slot;
    const uint8_t* __restrict__ pyrI = seq_base + (size_t)slot_i * lay.slot_bytes;
    const uint8_t* __restrict__ pyrJ = seq_base + (size_t)(1 - slot_i) * lay.slot_bytes;

    const float2 pt =
        in_pts[(size_t)cd.seq * in_stride + (in_buffered ? (size_t)cd.feat_buf * max_pts : 0) + pi];

    // window pixels owned by this thread
    int wofs[NPT];                           // wy * 32 + wx, or -1
#pragma unroll
    for (int t = 0; t < NPT; ++t) {
        int k = tid + t * NT;
        wofs[t] = (k < WIN2) ? ((k / WIN) << 5 | (k % WIN)) : -1;
    }

    const float half = 10.f;                 // (winSize - 1) * 0.5
    const float FLT_SCALE = 1.f / (float)(1 << 20);
    float2 nextPt = make_float2(0.f, 0.f);   // the value OpenCV keeps in nextPts[ptidx]
    bool ok = true;
    int red_par = 0;

    for (int l = lay.n_levels - 1; l >= 0; --l) {
        const SvoLevel lv = lay.lvl[l];
        const uint8_t* __restrict__ I = pyrI + lv.off;
        const uint8_t* __restrict__ J = pyrJ + lv.off;
        const int w = lv.w, h = lv.h, pitch = lv.pitch;
        const float inv = 1.f / (float)(1 << l);
        float2 prevPt = make_float2(__fmul_rn(pt.x, inv), __fmul_rn(pt.y, inv));
        if (l == lay.n_levels - 1) nextPt = prevPt;
        else nextPt = make_float2(__fmul_rn(nextPt.x, 2.f), __fmul_rn(nextPt.y, 2.f));

        prevPt.x = __fsub_rn(prevPt.x, half);
        prevPt.y = __fsub_rn(prevPt.y, half);
        const int ipx = (int)floorf(prevPt.x), ipy = (int)floorf(prevPt.y);
        if (ipx < -WIN || ipx >= w || ipy < -WIN || ipy >= h) {
            if (l == 0) ok = false;
            continue;
        }
        int w00, w01, w10, w11;
        lk_weights(__fsub_rn(prevPt.x, (float)ipx), __fsub_rn(prevPt.y, (float)ipy), w00, w01,
                   w10, w11);

        __syncthreads();                     // previous level's readers are done with smem
        for (int i = tid; i < IP * IP; i += NT) {
            int r = i / IP, c = i - r * IP;
            int gy = clamp_reflect(ipy - 1 + r, h), gx = clamp_reflect(ipx - 1 + c, w);
            s_I[r][c] = I[(size_t)gy * pitch + gx];
        }
        __syncthreads();
        for (int i = tid; i < DG * DG; i += NT) {
            int r = i / DG, c = i - r * DG;
            int gy = ipy + r, gx = ipx + c;
            int vx = 0, vy = 0;
            if (gx >= 0 && gx < w && gy >= 0 && gy < h) {
                // s_I[r + 1][c + 1] is pixel (gy, gx)
                int a0 = s_I[r][c], a1 = s_I[r][c + 1], a2 = s_I[r][c + 2];
                int b0 = s_I[r + 1][c], b2 = s_I[r + 1][c + 2];
                int c0 = s_I[r + 2][c], c1 = s_I[r + 2][c + 1], c2 = s_I[r + 2][c + 2];
                int t0l = 3 * (a0 + c0) + 10 * b0, t0r = 3 * (a2 + c2) + 10 * b2;
                int t1l = c0 - a0, t1c = c1 - a1, t1r = c2 - a2;
                vx = t0r - t0l;
                vy = 3 * (t1l + t1r) + 10 * t1c;
            }
            s_dx[r][c] = (short)vx;
            s_dy[r][c] = (short)vy;
        }
        __syncthreads();

        int ival[NPT], ixv[NPT], iyv[NPT];
        int sa11 = 0, sa12 = 0, sa22 = 0;
#pragma unroll
        for (int t = 0; t < NPT; ++t) {
            ival[t] = ixv[t] = iyv[t] = 0;
            if (wofs[t] >= 0) {
                int wy = wofs[t] >> 5, wx = wofs[t] & 31;
                ival[t] = (s_I[wy + 1][wx + 1] * w00 + s_I[wy + 1][wx + 2] * w01 +
                           s_I[wy + 2][wx + 1] * w10 + s_I[wy + 2][wx + 2] * w11 + 256) >> 9;
                ixv[t] = (s_dx[wy][wx] * w00 + s_dx[wy][wx + 1] * w01 + s_dx[wy + 1][wx] * w10 +
                          s_dx[wy + 1][wx + 1] * w11 + 8192) >> 14;
                iyv[t] = (s_dy[wy][wx] * w00 + s_dy[wy][wx + 1] * w01 + s_dy[wy + 1][wx] * w10 +
                          s_dy[wy + 1][wx + 1] * w11 + 8192) >> 14;
                sa11 += ixv[t] * ixv[t];
                sa12 += ixv[t] * iyv[t];
                sa22 += iyv[t] * iyv[t];
            }
        }
        long long iA11 = warp_sum_exact(sa11), iA12 = warp_sum_exact(sa12),
                  iA22 = warp_sum_exact(sa22);
        if (NW > 1) {
            if ((tid & 31) == 0) {
                s_red[red_par][wrp][0] = iA11;
                s_red[red_par][wrp][1] = iA12;
                s_red[red_par][wrp][2] = iA22;
            }
            __syncthreads();
            iA11 = iA12 = iA22 = 0;
#pragma unroll
            for (int q = 0; q < NW; ++q) {
                iA11 += s_red[red_par][q][0];
                iA12 += s_red[red_par][q][1];
                iA22 += s_red[red_par][q][2];
            }
            red_par ^= 1;
        }
        const float A11 = __fmul_rn(__ll2float_rn(iA11), FLT_SCALE);
        const float A12 = __fmul_rn(__ll2float_rn(iA12), FLT_SCALE);
        const float A22 = __fmul_rn(__ll2float_rn(iA22), FLT_SCALE);
        float D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float dif = __fsub_rn(A11, A22);
        const float rad = __fadd_rn(__fmul_rn(dif, dif), __fmul_rn(__fmul_rn(4.f, A12), A12));
        const float minEig =
            __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11), __fsqrt_rn(rad)), (float)(2 * WIN * WIN));
        if ((double)minEig < min_eig_thr || D < 1.1920928955078125e-7f) {
            if (l == 0) ok = false;
            continue;
        }
        D = __fdiv_rn(1.f, D);

        float2 np = make_float2(__fsub_rn(nextPt.x, half), __fsub_rn(nextPt.y, half));
        float2 prevDelta = make_float2(0.f, 0.f);
        int jx0 = -(1 << 30), jy0 = -(1 << 30);      // staged J origin (none yet)
        for (int j = 0; j < max_iters; ++j) {
            const int inx = (int)floorf(np.x), iny = (int)floorf(np.y);
            if (inx < -WIN || inx >= w || iny < -WIN || iny >= h) {
                if (l == 0) ok = false;
                break;
            }
            lk_weights(__fsub_rn(np.x, (float)inx), __fsub_rn(np.y, (float)iny), w00, w01, w10,
                       w11);
            int ox = inx - jx0, oy = iny - jy0;
            if (ox < 0 || ox > 2 * JM || oy < 0 || oy > 2 * JM) {   // uniform across the block
                __syncthreads();
                jx0 = inx - JM;
                jy0 = iny - JM;
                for (int i = tid; i < JR * JR; i += NT) {
                    int r = i / JR, c = i - r * JR;
                    int gy = clamp_reflect(jy0 + r, h), gx = clamp_reflect(jx0 + c, w);
                    s_J[r][c] = J[(size_t)gy * pitch + gx];
                }
                __syncthreads();
                ox = oy = JM;
            }
            int sb1 = 0, sb2 = 0;
#pragma unroll
            for (int t = 0; t < NPT; ++t) {
                if (wofs[t] >= 0) {
                    int wy = (wofs[t] >> 5) + oy, wx = (wofs[t] & 31) + ox;
                    int jv = (s_J[wy][wx] * w00 + s_J[wy][wx + 1] * w01 + s_J[wy + 1][wx] * w10 +
                              s_J[wy + 1][wx + 1] * w11 + 256) >> 9;
                    int diff = jv - ival[t];
                    sb1 += diff * ixv[t];
                    sb2 += diff * iyv[t];
                }
            }
            long long ib1 = warp_sum_exact(sb1), ib2 = warp_sum_exact(sb2);
            if (NW > 1) {
                if ((tid & 31) == 0) {
                    s_red[red_par][wrp][0] = ib1;
                    s_red[red_par][wrp][1] = ib2;
                }
                __syncthreads();
                ib1 = ib2 = 0;
#pragma unroll
                for (int q = 0; q < NW; ++q) {
                    ib1 += s_red[red_par][q][0];
                    ib2 += s_red[red_par][q][1];
                }
                red_par ^= 1;
            }
            const float b1 = __fmul_rn(__ll2float_rn(ib1), FLT_SCALE);
            const float b2 = __fmul_rn(__ll2float_rn(ib2), FLT_SCALE);
            float2 delta;
            delta.x = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), D);
            delta.y = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), D);
            np.x = __fadd_rn(np.x, delta.x);
            np.y = __fadd_rn(np.y, delta.y);
            nextPt = make_float2(__fadd_rn(np.x, half), __fadd_rn(np.y, half));
            const double dd = (double)delta.x * (double)delta.x + (double)delta.y * (double)delta.y;
            if (dd <= eps2) break;
            if (j > 0 && fabs((double)__fadd_rn(delta.x, prevDelta.x)) < 0.01 &&
                fabs((double)__fadd_rn(delta.y, prevDelta.y)) < 0.01) {
                nextPt.x = __fsub_rn(nextPt.x, __fmul_rn(delta.x, 0.5f));
                nextPt.y = __fsub_rn(nextPt.y, __fmul_rn(delta.y, 0.5f));
                break;
            }
            prevDelta = delta;
        }
    }
    // level-0 epilogue: `err` is requested by the reference (feature_tracker.cpp:24), so OpenCV
    // re-checks the final window origin and clears the status if it left the bordered image
    if (ok) {
        const SvoLevel l0 = lay.lvl[0];
        int inx = (int)floorf(__fsub_rn(nextPt.x, half)), iny = (int)floorf(__fsub_rn(nextPt.y, half));
        if (inx < -WIN || inx >= l0.w || iny < -WIN || iny >= l0.h) ok = false;
    }
    if (tid == 0) {
        out_pts[(size_t)cd.seq * out_stride + pi] = nextPt;
        status[(size_t)cd.seq * out_stride + pi] = ok ? 1 : 0;
    }
}

}  // namespace

cudaError_t svo_launch_lk_track(const uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                const float2* in_pts, size_t in_stride, int in_buffered,
                                float2* out_pts, unsigned char* status, int out_stride,
                                int direction, SvoLkParams lp, int max_pts, const int* gate,
                                int gate_min, const SvoSeqCall* d_call, int n_call,
                                cudaStream_t st) {
    if (max_pts <= 0 || n_call <= 0) return cudaSuccess;
    dim3 grid(lp.grid_pts > 0 ? lp.grid_pts : max_pts, n_call);
    if (lp.warps_per_point == 1)
        lk_track_kernel<1><<<grid, 32, 0, st>>>(pyr, lay, seq_stride, in_pts, in_stride,
                                                in_buffered, out_pts, status, out_stride,
                                                direction, lp.max_iters, lp.eps2, lp.min_eig,
                                                max_pts, gate, gate_min, d_call);
    else if (lp.warps_per_point == 4)
        lk_track_kernel<4><<<grid, 128, 0, st>>>(pyr, lay, seq_stride, in_pts, in_stride,
                                                 in_buffered, out_pts, status, out_stride,
                                                 direction, lp.max_iters, lp.eps2, lp.min_eig,
                                                 max_pts, gate, gate_min, d_call);
    else
        lk_track_kernel<2><<<grid, 64, 0, st>>>(pyr, lay, seq_stride, in_pts, in_stride,
                                                in_buffered, out_pts, status, out_stride,
                                                direction, lp.max_iters, lp.eps2, lp.min_eig,
                                                max_pts, gate, gate_min, d_call);
    return cudaGetLastError();
}
