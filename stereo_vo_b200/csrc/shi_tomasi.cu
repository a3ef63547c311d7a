// shi_tomasi_score: the min-eigenvalue map that cv::goodFeaturesToTrack computes first
// (reference call site src/image_processor.cpp:22; arithmetic pinned bit-for-bit in
// oracle/restate.py:min_eigen_map).
//
//   Dx = fl(fl(d[y]*k0) + fl(fl(d[y-1]+d[y+1])*k1)),  d = I[x+1]-I[x-1]
//   Dy = fl(s[y+1]-s[y-1]),  s = fl(fl(fl(I[x-1]*k1)+fl(I[x]*k0))+fl(I[x+1]*k1))
//   xx,xy,yy = fl(products); 3x3 box in fp64: row sums (c[x-1]+c[x])+c[x+1], column sums by a
//   SLIDING accumulator down the whole image (history-dependent rounding, so a column must be
//   walked top to bottom by one thread); eig = fl(fl(a+c) - sqrt(fl(fl((a-c)^2)+fl(b*b)))).
//
// Mapping: one warp owns 30 output columns (lanes 0 and 31 are halo columns) and walks down
// all rows; neighbours are exchanged with warp shuffles, so there is no shared memory and no
// block barrier. Every float op is an explicit round-to-nearest intrinsic (no FMA contraction).
// The global maximum (needed for the quality threshold) is folded in with one atomicMax per
// warp on a monotone integer key.
// Algorithmic bytes: w*h (u8 read) + 4*w*h (f32 write) = 5N per image.
#include "svo_common.cuh"

namespace {

constexpr int ST_WARPS = 2;
constexpr unsigned FULL = 0xffffffffu;

struct Chain {
    double ra, rb, sum;
};

__device__ __forceinline__ float eig_from_cov(float A, float B, float C) {
    float a = __fmul_rn(A, 0.5f);
    float c = __fmul_rn(C, 0.5f);
    float amc = __fsub_rn(a, c);
    float t = __fadd_rn(__fmul_rn(amc, amc), __fmul_rn(B, B));
    return __fsub_rn(__fadd_rn(a, c), __fsqrt_rn(t));
}

__global__ void __launch_bounds__(ST_WARPS * 32)
shi_tomasi_kernel(const uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride,
                  float* __restrict__ eig, int eig_pitch, size_t eig_stride,
                  unsigned int* __restrict__ eig_max, const SvoSeqCall* __restrict__ call) {
    const SvoSeqCall cd = call[blockIdx.y];
    const SvoLevel l0 = lay.lvl[0];
    const int W = l0.w, H = l0.h, pitch = l0.pitch;
    const uint8_t* __restrict__ img =
        pyr + (size_t)cd.seq * seq_stride + (size_t)cd.cur_slot * lay.slot_bytes + l0.off;
    float* __restrict__ out = eig + (size_t)cd.seq * eig_stride;

    const int lane = threadIdx.x & 31;
    const int warp = blockIdx.x * ST_WARPS + (threadIdx.x >> 5);
    const int xw0 = warp * 30;
    if (xw0 >= W) return;                      // whole warp exits together
    const int x = xw0 + lane - 1;
    const int xc = min(max(svo_reflect101(x, W), 0), W - 1);
    const bool has_extra = (lane == 0) || (lane == 31);
    const int xe = min(max(svo_reflect101(lane == 0 ? x - 1 : x + 1, W), 0), W - 1);
    const bool store_lane = (lane >= 1) && (lane <= 30) && (x < W);

    const float k1 = (float)(1.0 / (4.0 * 3.0 * 255.0));
    const float k0 = (float)(2.0 * (1.0 / (4.0 * 3.0 * 255.0)));

    auto load_row = [&](int y) -> unsigned {
        int yr = min(max(svo_reflect101(y, H), 0), H - 1);
        const uint8_t* rp = img + (size_t)yr * pitch;
        unsigned v = rp[xc];
        if (has_extra) v |= (unsigned)rp[xe] << 8;
        return v;
    };
    // (d, s) of one pixel row for this lane's column
    auto row_terms = [&](unsigned raw, float& d, float& s) {
        float P = (float)(raw & 0xffu);
        float E = (float)(raw >> 8);
        float L = __shfl_up_sync(FULL, P, 1);
        float R = __shfl_down_sync(FULL, P, 1);
        if (lane == 0) L = E;
        if (lane == 31) R = E;
        d = __fsub_rn(R, L);
        s = __fadd_rn(__fadd_rn(__fmul_rn(L, k1), __fmul_rn(P, k0)), __fmul_rn(R, k1));
    };

    float d0, s0, d1, s1;                      // rows j-1 and j
    row_terms(load_row(-1), d0, s0);
    row_terms(load_row(0), d1, s1);

    Chain cxx_c = {0, 0, 0}, cxy_c = {0, 0, 0}, cyy_c = {0, 0, 0};
    float vmax = -__int_as_float(0x7f800000);  // -inf

    auto emit = [&](int y, double sxx, double sxy, double syy) {
        float e = eig_from_cov((float)sxx, (float)sxy, (float)syy);
        if (store_lane) {
            out[(size_t)y * eig_pitch + x] = e;
            vmax = fmaxf(vmax, e);
        }
    };

    // prefetch queue: pixel rows j+1 .. j+4
    unsigned q0 = load_row(1), q1 = load_row(2), q2 = load_row(3), q3 = load_row(4);

    auto step = [&](unsigned raw, int j) {
        float dn, sn;
        row_terms(raw, dn, sn);                // pixel row j+1
        float Dx = __fadd_rn(__fmul_rn(d1, k0), __fmul_rn(__fadd_rn(d0, dn), k1));
        float Dy = __fsub_rn(sn, s0);
        d0 = d1; d1 = dn; s0 = s1; s1 = sn;
        float cxx = __fmul_rn(Dx, Dx), cxy = __fmul_rn(Dx, Dy), cyy = __fmul_rn(Dy, Dy);
        float lxx = __shfl_up_sync(FULL, cxx, 1), rxx = __shfl_down_sync(FULL, cxx, 1);
        float lxy = __shfl_up_sync(FULL, cxy, 1), rxy = __shfl_down_sync(FULL, cxy, 1);
        float lyy = __shfl_up_sync(FULL, cyy, 1), ryy = __shfl_down_sync(FULL, cyy, 1);
        // box-filter border is reflect-101 on the PRODUCT images: c[-1]=c[1], c[W]=c[W-2]
        if (x == 0) { lxx = rxx; lxy = rxy; lyy = ryy; }
        if (x == W - 1) { rxx = lxx; rxy = lxy; ryy = lyy; }
        double r_xx = __dadd_rn(__dadd_rn((double)lxx, (double)cxx), (double)rxx);
        double r_xy = __dadd_rn(__dadd_rn((double)lxy, (double)cxy), (double)rxy);
        double r_yy = __dadd_rn(__dadd_rn((double)lyy, (double)cyy), (double)ryy);
        if (j == 0) {
            cxx_c.rb = r_xx; cxy_c.rb = r_xy; cyy_c.rb = r_yy;
            return;
        }
        if (j == 1) {                          // SUM = (0 + r[-1]) + r[0], r[-1] = r[1]
            cxx_c.sum = __dadd_rn(__dadd_rn(0.0, r_xx), cxx_c.rb); cxx_c.ra = r_xx;
            cxy_c.sum = __dadd_rn(__dadd_rn(0.0, r_xy), cxy_c.rb); cxy_c.ra = r_xy;
            cyy_c.sum = __dadd_rn(__dadd_rn(0.0, r_yy), cyy_c.rb); cyy_c.ra = r_yy;
        }
        double sxx = __dadd_rn(cxx_c.sum, r_xx);
        double sxy = __dadd_rn(cxy_c.sum, r_xy);
        double syy = __dadd_rn(cyy_c.sum, r_yy);
        emit(j - 1, sxx, sxy, syy);
        cxx_c.sum = __dsub_rn(sxx, cxx_c.ra); cxx_c.ra = cxx_c.rb; cxx_c.rb = r_xx;
        cxy_c.sum = __dsub_rn(sxy, cxy_c.ra); cxy_c.ra = cxy_c.rb; cxy_c.rb = r_xy;
        cyy_c.sum = __dsub_rn(syy, cyy_c.ra); cyy_c.ra = cyy_c.rb; cyy_c.rb = r_yy;
    };

    for (int j = 0; j < H; j += 4) {
        step(q0, j);
        q0 = load_row(j + 5);
        if (j + 1 < H) step(q1, j + 1);
        q1 = load_row(j + 6);
        if (j + 2 < H) step(q2, j + 2);
        q2 = load_row(j + 7);
        if (j + 3 < H) step(q3, j + 3);
        q3 = load_row(j + 8);
    }
    // last row: r[H] = r[H-2], which is `ra` after the final shift
    emit(H - 1, __dadd_rn(cxx_c.sum, cxx_c.ra), __dadd_rn(cxy_c.sum, cxy_c.ra),
         __dadd_rn(cyy_c.sum, cyy_c.ra));

    unsigned key = store_lane ? svo_float_key(__float_as_uint(vmax)) : 0u;
    key = __reduce_max_sync(FULL, key);
    if (lane == 0) atomicMax(eig_max + cd.seq, key);
}

}  // namespace

cudaError_t svo_launch_shi_tomasi(const uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                  float* eig, int eig_pitch, size_t eig_stride,
                                  unsigned int* eig_max, const SvoSeqCall* d_call, int n_call,
                                  cudaStream_t st) {
    int warps = (lay.lvl[0].w + 29) / 30;
    dim3 grid((warps + ST_WARPS - 1) / ST_WARPS, n_call);
    shi_tomasi_kernel<<<grid, ST_WARPS * 32, 0, st>>>(pyr, lay, seq_stride, eig, eig_pitch,
                                                      eig_stride, eig_max, d_call);
    return cudaGetLastError();
}
