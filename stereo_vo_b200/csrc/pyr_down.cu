// pyr_down_u8: one level of the 8-bit Gaussian pyramid that cv::calcOpticalFlowPyrLK builds
// internally (reference call sites src/feature_tracker.cpp:23-26,32-35; arithmetic pinned in
// oracle/restate.py:pyr_down). dst size ((w+1)/2, (h+1)/2); separable [1 4 6 4 1] with
// BORDER_REFLECT_101; out = (sum + 128) >> 8. Pure integer, bit-exact.
//
// HBM-bound: algorithmic bytes = w*h (read) + ow*oh (write). A block produces a 128x8 output
// tile: the (2*8+3) x (256+32) source footprint is staged into shared memory with 128-bit
// loads (rows reflected at staging time, columns reflected when the taps are read), the
// horizontal pass leaves 16-bit partial sums in shared memory, the vertical pass writes packed
// uchar4 words (a warp stores 128 contiguous bytes).
#include "svo_common.cuh"

namespace {

constexpr int TW = 128;                  // output tile width
constexpr int TH = 8;                    // output tile height
constexpr int SROWS = 2 * TH + 3;        // source rows needed: 2*oy-2 .. 2*(oy+TH-1)+2
constexpr int SCOLS = 2 * TW + 32;       // staged source columns: [2*ox0-16, 2*ox0+2*TW+16)
constexpr int NTHREADS = 256;

__device__ __forceinline__ void pyr_down_tile(const uint8_t* __restrict__ src, int w, int h,
                                              int spitch, uint8_t* __restrict__ dst, int ow,
                                              int oh, int dpitch, int ox0, int oy0) {
    __shared__ __align__(16) uint8_t s_src[SROWS][SCOLS];
    __shared__ __align__(16) unsigned short s_h[SROWS][TW];

    const int tid = threadIdx.x;
    const int sx0 = 2 * ox0 - 16;        // multiple of 16 (ox0 % 128 == 0)
    const int sy0 = 2 * oy0 - 2;
    // ---- stage: SROWS x (SCOLS/16) chunks of 16 bytes
    constexpr int CHUNKS = SCOLS / 16;
    for (int i = tid; i < SROWS * CHUNKS; i += NTHREADS) {
        int r = i / CHUNKS, c = i - r * CHUNKS;
        int gy = svo_reflect101(sy0 + r, h);
        gy = min(max(gy, 0), h - 1);         // rows feeding outputs >= oh only (never stored)
        int gx = sx0 + c * 16;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (gx >= 0 && gx + 16 <= spitch)   // pitch is a multiple of 16; pad bytes are never used
            v = *reinterpret_cast<const uint4*>(src + (size_t)gy * spitch + gx);
        *reinterpret_cast<uint4*>(&s_src[r][c * 16]) = v;
    }
    __syncthreads();
    // ---- horizontal pass: s_h[r][ox] = I[2x-2] + 4 I[2x-1] + 6 I[2x] + 4 I[2x+1] + I[2x+2]
    for (int i = tid; i < SROWS * TW; i += NTHREADS) {
        int r = i / TW, ox = i - r * TW;
        int gx = 2 * (ox0 + ox);
        int v = 0;
        if (ox0 + ox < ow) {
            int x0 = svo_reflect101(gx - 2, w) - sx0;
            int x1 = svo_reflect101(gx - 1, w) - sx0;
            int x2 = gx - sx0;
            int x3 = svo_reflect101(gx + 1, w) - sx0;
            int x4 = svo_reflect101(gx + 2, w) - sx0;
            const uint8_t* row = s_src[r];
            v = row[x0] + row[x4] + 4 * (row[x1] + row[x3]) + 6 * row[x2];
        }
        s_h[r][ox] = (unsigned short)v;
    }
    __syncthreads();
    // ---- vertical pass: 4 outputs per thread, one uchar4 store
    {
        int oyl = tid / (TW / 4);            // 0..7
        int oxl = (tid - oyl * (TW / 4)) * 4;
        int oy = oy0 + oyl;
        if (oy < oh) {
            unsigned int packed = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int r = 2 * oyl;
                int v = s_h[r][oxl + k] + s_h[r + 4][oxl + k] +
                        4 * (s_h[r + 1][oxl + k] + s_h[r + 3][oxl + k]) + 6 * s_h[r + 2][oxl + k];
                packed |= (unsigned int)((v + 128) >> 8) << (8 * k);
            }
            int ox = ox0 + oxl;
            if (ox < ow)                      // dpitch % 4 == 0, so a whole word always fits
                *reinterpret_cast<unsigned int*>(dst + (size_t)oy * dpitch + ox) = packed;
        }
    }
}

__global__ void __launch_bounds__(NTHREADS)
pyr_down_u8_kernel(uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride, int level,
                   const SvoSeqCall* __restrict__ call) {
    const SvoSeqCall c = call[blockIdx.z];
    uint8_t* slot = pyr + (size_t)c.seq * seq_stride + (size_t)c.cur_slot * lay.slot_bytes;
    const SvoLevel s = lay.lvl[level];
    const SvoLevel d = lay.lvl[level + 1];
    pyr_down_tile(slot + s.off, s.w, s.h, s.pitch, slot + d.off, d.w, d.h, d.pitch,
                  blockIdx.x * TW, blockIdx.y * TH);
}

__global__ void __launch_bounds__(NTHREADS)
pyr_down_u8_single_kernel(const uint8_t* __restrict__ src, int w, int h, int spitch,
                          uint8_t* __restrict__ dst, int dpitch) {
    pyr_down_tile(src, w, h, spitch, dst, (w + 1) / 2, (h + 1) / 2, dpitch, blockIdx.x * TW,
                  blockIdx.y * TH);
}

}  // namespace

cudaError_t svo_launch_pyr_down(uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    for (int l = 0; l + 1 < lay.n_levels; ++l) {
        const SvoLevel& d = lay.lvl[l + 1];
        dim3 grid((d.w + TW - 1) / TW, (d.h + TH - 1) / TH, n_call);
        pyr_down_u8_kernel<<<grid, NTHREADS, 0, st>>>(pyr, lay, seq_stride, l, d_call);
    }
    return cudaGetLastError();
}

cudaError_t svo_launch_pyr_down_single(const uint8_t* src, int w, int h, int spitch,
                                       uint8_t* dst, int dpitch, cudaStream_t st) {
    int ow = (w + 1) / 2, oh = (h + 1) / 2;
    dim3 grid((ow + TW - 1) / TW, (oh + TH - 1) / TH, 1);
    pyr_down_u8_single_kernel<<<grid, NTHREADS, 0, st>>>(src, w, h, spitch, dst, dpitch);
    return cudaGetLastError();
}
