// nms_threshold_compact: steps 6-7 of cv::goodFeaturesToTrack (reference call site
// src/image_processor.cpp:22; pinned in oracle/restate.py:corner_candidates).
//   thr = (float)((double)max(eig) * qualityLevel);  e' = eig > thr ? eig : 0
//   candidate <=> interior pixel, e' != 0 and e' == max of e' over its 3x3 neighbourhood
// Output: unordered list of 64-bit keys (monotone score key << 32 | y*W + x); the sort that
// follows orders them by (score desc, address desc), which is a total order, so the result is
// deterministic no matter in which order warps append.
//
// HBM-bound: algorithmic bytes = 4*w*h (one read of the f32 map) + 8 B per candidate.
// A lane reads one float4 (128-bit) per row, a warp 512 contiguous bytes; the horizontal
// 3-max uses shuffles, the vertical 3-max a register window while walking down a strip of rows,
// appends are warp-aggregated (ballot + one atomicAdd per warp and row).
#include "svo_common.cuh"

namespace {

constexpr int NMS_WARPS = 4;
constexpr int NMS_ROWS = 16;               // rows per warp strip
constexpr unsigned FULL = 0xffffffffu;

__global__ void __launch_bounds__(NMS_WARPS * 32)
nms_compact_kernel(const float* __restrict__ eig, int W, int H, int eig_pitch, size_t eig_stride,
                   const unsigned int* __restrict__ eig_max, double quality,
                   unsigned long long* __restrict__ cand, int max_cand,
                   int* __restrict__ cand_count, const SvoSeqCall* __restrict__ call) {
    const int seq = call[blockIdx.z].seq;
    const float* __restrict__ e = eig + (size_t)seq * eig_stride;
    unsigned long long* __restrict__ out = cand + (size_t)seq * max_cand;
    int* __restrict__ cnt = cand_count + seq;

    const int lane = threadIdx.x & 31;
    const int wrp = threadIdx.x >> 5;
    const int x4 = (blockIdx.x * 32 + lane) * 4;
    const int ys = (blockIdx.y * NMS_WARPS + wrp) * NMS_ROWS;
    if (ys >= H) return;
    const float maxv = __uint_as_float(svo_key_float_bits(eig_max[seq]));
    const float thr = (float)((double)maxv * quality);
    const float NEG = -__int_as_float(0x7f800000);
    const bool in_pitch = x4 < eig_pitch;    // pitch % 4 == 0

    float4 ec = make_float4(0, 0, 0, 0);     // e' of the centre row (y-1)
    float4 h0, h1;                            // horizontal 3-max of rows y-2, y-1
    h0 = h1 = make_float4(NEG, NEG, NEG, NEG);

    const int yend = min(ys + NMS_ROWS, H - 1);   // candidates live in rows [max(ys,1), yend)
    for (int y = ys - 1; y <= yend; ++y) {
        float4 v = make_float4(NEG, NEG, NEG, NEG);
        float lft = NEG, rgt = NEG;
        if (y >= 0 && y < H && in_pitch) {
            const float* row = e + (size_t)y * eig_pitch;
            v = *reinterpret_cast<const float4*>(row + x4);
            if (lane == 0 && x4 > 0) lft = row[x4 - 1];
            if (lane == 31 && x4 + 4 < eig_pitch) rgt = row[x4 + 4];
            v.x = v.x > thr ? v.x : 0.f;
            v.y = v.y > thr ? v.y : 0.f;
            v.z = v.z > thr ? v.z : 0.f;
            v.w = v.w > thr ? v.w : 0.f;
            lft = lft > thr ? lft : 0.f;
            rgt = rgt > thr ? rgt : 0.f;
        }
        float fl = __shfl_up_sync(FULL, v.w, 1);
        float fr = __shfl_down_sync(FULL, v.x, 1);
        if (lane != 0) lft = fl;
        if (lane != 31) rgt = fr;
        float4 h2;
        h2.x = fmaxf(fmaxf(lft, v.x), v.y);
        h2.y = fmaxf(fmaxf(v.x, v.y), v.z);
        h2.z = fmaxf(fmaxf(v.y, v.z), v.w);
        h2.w = fmaxf(fmaxf(v.z, v.w), rgt);
        const int yc = y - 1;                 // row being decided
        if (yc >= ys && yc >= 1 && yc < yend) {
            float m[4] = {fmaxf(fmaxf(h0.x, h1.x), h2.x), fmaxf(fmaxf(h0.y, h1.y), h2.y),
                          fmaxf(fmaxf(h0.z, h1.z), h2.z), fmaxf(fmaxf(h0.w, h1.w), h2.w)};
            float c[4] = {ec.x, ec.y, ec.z, ec.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int x = x4 + k;
                bool is = (x >= 1) && (x <= W - 2) && (c[k] != 0.f) && (c[k] == m[k]);
                unsigned bal = __ballot_sync(FULL, is);
                if (bal) {
                    int base = 0;
                    if (lane == 0) base = atomicAdd(cnt, __popc(bal));
                    base = __shfl_sync(FULL, base, 0);
                    if (is) {
                        int idx = base + __popc(bal & ((1u << lane) - 1u));
                        if (idx < max_cand) {
                            unsigned long long key =
                                ((unsigned long long)svo_float_key(__float_as_uint(c[k])) << 32) |
                                (unsigned int)(yc * W + x);
                            out[idx] = key;
                        }
                    }
                }
            }
        }
        h0 = h1; h1 = h2; ec = v;
    }
}

}  // namespace

cudaError_t svo_launch_nms_compact(const float* eig, int w, int h, int eig_pitch,
                                   size_t eig_stride, const unsigned int* eig_max,
                                   double quality, unsigned long long* cand, int max_cand,
                                   int* cand_count, const SvoSeqCall* d_call, int n_call,
                                   cudaStream_t st) {
    dim3 grid((eig_pitch / 4 + 31) / 32, (h + NMS_WARPS * NMS_ROWS - 1) / (NMS_WARPS * NMS_ROWS),
              n_call);
    nms_compact_kernel<<<grid, NMS_WARPS * 32, 0, st>>>(eig, w, h, eig_pitch, eig_stride, eig_max,
                                                        quality, cand, max_cand, cand_count,
                                                        d_call);
    return cudaGetLastError();
}
