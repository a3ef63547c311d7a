// pyr_down_u8: one level of the 8-bit Gaussian pyramid that cv::calcOpticalFlowPyrLK builds
// internally (reference call sites src/feature_tracker.cpp:23-26,32-35; arithmetic pinned in
// oracle/restate.py:pyr_down). dst size ((w+1)/2, (h+1)/2); separable [1 4 6 4 1] with
// BORDER_REFLECT_101; out = (sum + 128) >> 8. Pure integer, bit-exact.
//
// HBM-bound: algorithmic bytes = w*h (read) + ow*oh (write). One launch per level, chained by
// programmatic dependent launch (the grid of level l+1 is scheduled while level l still runs
// and waits for its stores in cudaGridDependencySynchronize): levels 2 and 3 are 2.3 and 0.6 MB
// of work for 64 KITTI images - their cost was launch latency. A single launch that hands out
// the tiles of all levels through an atomic counter and per-image completion flags was built
// and measured slower (44 us against 27 us for three plain launches and 23 us with dependent
// launch: every block paid a ticket, a fence and two atomics). Register-only tiles, no shared
// memory:
//   * a thread owns 4 adjacent output columns and ROWS output rows; a warp therefore
//     reads 256 contiguous source bytes per source row (aligned 32/64-bit loads) and stores 128
//     contiguous output bytes per output row;
//   * horizontal pass: the five taps of an output are one dp4a with weights (1,4,6,4) on the
//     word that starts two pixels left of the centre plus one dp4a that picks the fifth byte -
//     8 dp4a + 2 byte permutes per 4 outputs and source row;
//   * the 12-bit horizontal sums of two neighbouring outputs share one 32-bit register (16-bit
//     lanes); the vertical pass P0 + P4 + 4(P1 + P3) + 6 P2 + 128 cannot carry between lanes
//     (<= 255*256 + 128 < 2^16), so it runs on both lanes at once;
//   * a thread filters its 2*ROWS+3 source rows up front (all loads in flight together), then
//     runs the vertical pass; rows outside the image are fetched through reflect-101 indices,
//     columns outside the image by per-thread byte-permute / weight constants (no edge path).
#include "svo_common.cuh"

namespace {

constexpr int BANDS = 4;                 // row bands per block (threadIdx.y)
constexpr unsigned W_1464 = 0x04060401u; // byte0..3 weights 1,4,6,4

// Per-thread constants of the horizontal pass. Column reflect-101 never needs a second code
// path: on the left only output 0 reads columns -2,-1 (= columns 2,1: a different byte
// permute of the same word); on the right only the LAST output column has taps outside the
// image (centre c = w-1 for odd w: taps c+1,c+2 fold onto c-1,c-2; c = w-2 for even w: tap c+2
// folds onto c), which is a different weight vector for that one output.
struct ColSetup {
    unsigned wgt[4];                     // dp4a weights of bytes c-2 .. c+1 per output
    unsigned fifth[4];                   // dp4a weights that pick byte c+2 (0: folded / unused)
    unsigned sel0;                       // byte permute building bytes x0-2 .. x0+1
    int off_m;                           // byte offset of the word left of x0 (clamped at 0)
};

__device__ __forceinline__ ColSetup col_setup(int w, int ow, int ox) {
    ColSetup cs;
    const unsigned f[4] = {0x00010000u, 0x00000001u, 0x00010000u, 0x00000001u};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const bool last = (ox + k == ow - 1);
        cs.wgt[k] = !last ? W_1464 : ((w & 1) ? 0x00060802u : 0x04070401u);
        cs.fifth[k] = !last ? f[k] : 0u;
    }
    cs.sel0 = (ox == 0) ? 0x5412u : 0x5432u;   // (I2, I1, I0, I1) at the left border
    cs.off_m = (ox == 0) ? 0 : 2 * ox - 4;
    return cs;
}

struct HRow {
    unsigned a, b;                       // (h0 | h1 << 16), (h2 | h3 << 16)
};

// horizontally filtered source row for output columns ox..ox+3 (ox % 4 == 0); `row` points at
// column 0 of the source row. Words may extend past column w-1 (into the row padding or the next
// row, always inside the allocation): those bytes only ever meet zero weights.
__device__ __forceinline__ HRow filter_row(const uint8_t* __restrict__ row, int x0,
                                           const ColSetup& cs) {
    const unsigned wm = __ldg(reinterpret_cast<const unsigned*>(row + cs.off_m));
    const uint2 c = __ldg(reinterpret_cast<const uint2*>(row + x0));
    const unsigned w2 = __ldg(reinterpret_cast<const unsigned*>(row + x0 + 8));
    const unsigned w0 = c.x, w1 = c.y;
    const unsigned m0 = __byte_perm(wm, w0, cs.sel0);  // bytes x0-2 .. x0+1 (off_m == 0: wm == w0)
    const unsigned m2 = __byte_perm(w0, w1, 0x5432);   // bytes x0+2 .. x0+5
    unsigned h0 = __dp4a(m0, cs.wgt[0], 0u);
    unsigned h1 = __dp4a(w0, cs.wgt[1], 0u);
    unsigned h2 = __dp4a(m2, cs.wgt[2], 0u);
    unsigned h3 = __dp4a(w1, cs.wgt[3], 0u);
    h0 = __dp4a(w0, cs.fifth[0], h0);                  // + I[x0+2]
    h1 = __dp4a(w1, cs.fifth[1], h1);                  // + I[x0+4]
    h2 = __dp4a(w1, cs.fifth[2], h2);                  // + I[x0+6]
    h3 = __dp4a(w2, cs.fifth[3], h3);                  // + I[x0+8]
    HRow r;
    r.a = h0 | (h1 << 16);
    r.b = h2 | (h3 << 16);
    return r;
}

__device__ __forceinline__ unsigned vertical(unsigned p0, unsigned p1, unsigned p2, unsigned p3,
                                             unsigned p4) {
    // both 16-bit lanes at once; no carry crosses bit 16 (see header)
    const unsigned v = p0 + p4 + 4u * (p1 + p3) + 6u * p2 + 0x00800080u;
    return (v >> 8) & 0x00ff00ffu;
}

// one thread's 4 output columns x ROWS output rows of level s -> level d
template <int ROWS>
__device__ __forceinline__ void pyr_tile(uint8_t* __restrict__ slot, const SvoLevel s,
                                         const SvoLevel d, int ox, int oy0) {
    if (ox >= d.w || oy0 >= d.h) return;
    const uint8_t* __restrict__ src = slot + s.off;
    uint8_t* __restrict__ dst = slot + d.off;
    const int h = s.h, spitch = s.pitch, x0 = 2 * ox;
    const ColSetup cs = col_setup(s.w, d.w, ox);
    auto srow = [&](int y) {                 // reflect-101 on rows, clamped for unused rows
        int yr = abs(y);
        yr = (yr >= h) ? 2 * h - 2 - yr : yr;
        yr = min(max(yr, 0), h - 1);
        return filter_row(src + (size_t)yr * spitch, x0, cs);
    };
    HRow r[2 * ROWS + 3];
#pragma unroll
    for (int i = 0; i < 2 * ROWS + 3; ++i) r[i] = srow(2 * oy0 - 2 + i);
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
        const int oy = oy0 + j;
        const unsigned lo = vertical(r[2 * j].a, r[2 * j + 1].a, r[2 * j + 2].a, r[2 * j + 3].a,
                                     r[2 * j + 4].a);
        const unsigned hi = vertical(r[2 * j].b, r[2 * j + 1].b, r[2 * j + 2].b, r[2 * j + 3].b,
                                     r[2 * j + 4].b);
        // dpitch % 4 == 0, so a whole word always fits (pad bytes are never read as pixels)
        if (oy < d.h)
            *reinterpret_cast<unsigned*>(dst + (size_t)oy * d.pitch + ox) =
                __byte_perm(lo, hi, 0x6420);
    }
}

template <int ROWS>
__global__ void __launch_bounds__(32 * BANDS)
pyr_down_u8_kernel(uint8_t* __restrict__ pyr, const __grid_constant__ SvoPyrLayout lay,
                   size_t seq_stride, int level, const SvoSeqCall* __restrict__ call,
                   unsigned int* __restrict__ img_serial, unsigned int serial) {
    // programmatic dependent launch: the next level's grid may be scheduled while this one
    // runs (its blocks wait below until this grid's stores are visible), so the three launches
    // of a pyramid overlap their launch latency and ramp-up instead of paying them in series
    cudaTriggerProgrammaticLaunchCompletion();
    const SvoSeqCall* c = call + blockIdx.z;
    uint8_t* slot = pyr + (size_t)c->seq * seq_stride + (size_t)c->cur_slot * lay.slot_bytes;
    if (level > 0) cudaGridDependencySynchronize();
    // a new image in this slot: whatever lk_track cached about the old one is void
    else if (img_serial && (blockIdx.x | blockIdx.y | threadIdx.x | threadIdx.y) == 0)
        img_serial[c->seq * 2 + c->cur_slot] = serial;
    pyr_tile<ROWS>(slot, lay.lvl[level], lay.lvl[level + 1],
                          (blockIdx.x * 32 + threadIdx.x) * 4,
                          (blockIdx.y * BANDS + threadIdx.y) * ROWS);
}

}  // namespace

cudaError_t svo_launch_pyr_down(uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                const SvoSeqCall* d_call, int n_call, cudaStream_t st,
                                unsigned int* img_serial, unsigned int serial) {
    // fewer rows per thread on the small levels: the grid, not the row overlap, is what counts
    for (int l = 0; l + 1 < lay.n_levels; ++l) {
        const SvoLevel& d = lay.lvl[l + 1];
        const int rows = (l == 0) ? 8 : (l == 1) ? 4 : 2;
        cudaLaunchConfig_t cfg = {};
        cfg.blockDim = dim3(32, BANDS);
        cfg.gridDim = dim3((d.w + 127) / 128, (d.h + rows * BANDS - 1) / (rows * BANDS), n_call);
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = l > 0 ? 1 : 0;       // levels 2, 3 may start under the level before
        cudaError_t ce;
        if (rows == 8) ce = cudaLaunchKernelEx(&cfg, pyr_down_u8_kernel<8>, pyr, lay, seq_stride, l, d_call, img_serial, serial);
        else if (rows == 4) ce = cudaLaunchKernelEx(&cfg, pyr_down_u8_kernel<4>, pyr, lay, seq_stride, l, d_call, img_serial, serial);
        else ce = cudaLaunchKernelEx(&cfg, pyr_down_u8_kernel<2>, pyr, lay, seq_stride, l, d_call, img_serial, serial);
        if (ce != cudaSuccess) return ce;
    }
    return cudaGetLastError();
}
