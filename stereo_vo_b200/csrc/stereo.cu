// Stereo stage of ImageProcessor::triangulate_stereo (reference src/image_processor.cpp:165-208)
// and the new-vs-tracked dedupe (src/image_processor.cpp:113-128).
//
// stereo_bm_sparse: cv::StereoBM::create(48, 21)->compute() evaluated ONLY at the requested
// pixels. The reference computes the dense map and then reads <= 300 pixels of it (:193); with
// the default parameters (no speckle filter, disp12MaxDiff = -1) a pixel's disparity depends
// only on its own 21-row x 68-column neighbourhood, so the sparse evaluation is bit-identical
// (arithmetic pinned in oracle/restate.py:stereo_bm_at):
//   XSOBEL prefilter (cap 31) of both images, 48 SADs over 21x21, first minimum in internal
//   order d = 47 - D, texture threshold 10, uniqueness ratio 15, parabola sub-pixel to 1/16 px.
// One block (64 threads) per key-point: raw patches -> shared memory, prefilter in shared
// memory, thread D accumulates SAD(D). Integer throughout. Issue/shared-memory bound.
//
// reproject_q: w = pose * Q * [x, y, d, 1]^T, point = w.xyz / w.w for d > 0, compacted in
// input order (:178-207). dedupe_new_features: keep detection i iff no inlier within
// min_feature_distance (f32 sqrt), ordered.
#include <stdlib.h>

#include "svo_common.cuh"

namespace {

constexpr int SB_THREADS = 64;
constexpr unsigned FULL = 0xffffffffu;

__global__ void __launch_bounds__(SB_THREADS)
stereo_bm_sparse_kernel(const uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride,
                        const uint8_t* __restrict__ right, size_t right_stride,
                        const float2* __restrict__ pts, int pts_stride,
                        short* __restrict__ disp16, SvoStereoParams sp,
                        const SvoSeqCall* __restrict__ call) {
    extern __shared__ unsigned char s_mem[];
    const SvoSeqCall cd = call[blockIdx.y];
    const int pi = blockIdx.x;
    if (pi >= cd.n_pts) return;
    const SvoLevel l0 = lay.lvl[0];
    const int W = l0.w, H = l0.h, pitch = l0.pitch;
    const uint8_t* __restrict__ L =
        pyr + (size_t)cd.seq * seq_stride + (size_t)cd.cur_slot * lay.slot_bytes + l0.off;
    const uint8_t* __restrict__ R = right + (size_t)cd.seq * right_stride;
    const float2 p = pts[(size_t)cd.seq * pts_stride + pi];
    short* outp = disp16 + (size_t)cd.seq * pts_stride + pi;
    const int x = (int)p.x, y = (int)p.y;     // float -> int truncation, image_processor.cpp:193

    const int nd = sp.ndisp, win = sp.win, rad = win / 2, cap = sp.cap;
    const short FILTERED = (short)(-16);      // (minDisparity - 1) << 4 with minDisparity = 0
    if (x < nd - 1 + rad || x >= W - rad || y < rad || y >= H - rad) {
        if (threadIdx.x == 0) *outp = FILTERED;
        return;
    }
    const int rw = win + nd - 1;              // prefiltered right-strip width
    const int rawLw = win + 2, rawRw = rw + 2, rawH = win + 2;
    // shared-memory carve-up
    unsigned char* rawL = s_mem;                              // rawH x rawLw
    unsigned char* rawR = rawL + rawH * rawLw;                // rawH x rawRw
    unsigned char* PL = rawR + rawH * rawRw;                  // win x win
    unsigned char* PR = PL + win * win;                       // win x rw
    int* s_sad = reinterpret_cast<int*>(s_mem + (((rawH * rawLw + rawH * rawRw + win * win +
                                                   win * rw) + 15) & ~15));   // nd + 2
    int* s_tex = s_sad + nd + 2;

    const int tid = threadIdx.x;
    const bool odd_last = (H & 1);
    // ---- stage raw patches (rows reflected the way prefilterXSobel walks them)
    const int yl0 = y - rad - 1;
    const int xl0 = x - rad - 1;
    const int xr0 = x - rad - (nd - 1) - 1;
    for (int i = tid; i < rawH * rawLw; i += SB_THREADS) {
        int r = i / rawLw, c = i - r * rawLw;
        int gy = yl0 + r;
        gy = gy < 0 ? 1 : (gy > H - 1 ? H - 2 : gy);
        int gx = min(max(xl0 + c, 0), W - 1);
        rawL[i] = L[(size_t)gy * pitch + gx];
    }
    for (int i = tid; i < rawH * rawRw; i += SB_THREADS) {
        int r = i / rawRw, c = i - r * rawRw;
        int gy = yl0 + r;
        gy = gy < 0 ? 1 : (gy > H - 1 ? H - 2 : gy);
        int gx = min(max(xr0 + c, 0), W - 1);
        rawR[i] = R[(size_t)gy * pitch + gx];
    }
    if (tid == 0) *s_tex = 0;
    __syncthreads();
    // ---- XSOBEL prefilter into PL / PR
    auto prefilter = [&](const unsigned char* raw, int raww, int r, int c, int gx, int gy) -> int {
        // (r, c) index the prefiltered window; raw is offset by one row / column
        if (gx == 0 || gx == W - 1 || (odd_last && gy == H - 1)) return cap;
        const unsigned char* a = raw + r * raww + c;          // row gy-1, col gx-1
        const unsigned char* b = a + raww;
        const unsigned char* d = b + raww;
        int v = ((int)a[2] - (int)a[0]) + 2 * ((int)b[2] - (int)b[0]) + ((int)d[2] - (int)d[0]);
        v = min(max(v, -cap), cap);
        return v + cap;
    };
    int tex = 0;
    for (int i = tid; i < win * win; i += SB_THREADS) {
        int r = i / win, c = i - r * win;
        int v = prefilter(rawL, rawLw, r, c, x - rad + c, y - rad + r);
        PL[i] = (unsigned char)v;
        tex += abs(v - cap);
    }
    for (int i = tid; i < win * rw; i += SB_THREADS) {
        int r = i / rw, c = i - r * rw;
        PR[i] = (unsigned char)prefilter(rawR, rawRw, r, c, x - rad - (nd - 1) + c, y - rad + r);
    }
    atomicAdd(s_tex, tex);
    __syncthreads();
    // ---- SAD(D) for D = tid, tid + 64, ...
    for (int D = tid; D < nd; D += SB_THREADS) {
        int sad = 0;
        const int shift = nd - 1 - D;
        for (int r = 0; r < win; ++r) {
            const unsigned char* pl = PL + r * win;
            const unsigned char* pr = PR + r * rw + shift;
#pragma unroll 7
            for (int c = 0; c < win; ++c) sad += abs((int)pl[c] - (int)pr[c]);
        }
        s_sad[1 + (nd - 1 - D)] = sad;        // internal order d = nd-1-D
    }
    __syncthreads();
    if (tid != 0) return;
    // ---- winner, texture, uniqueness, sub-pixel (findStereoCorrespondenceBM tail)
    int* sad = s_sad + 1;
    int mind = 0, minsad = sad[0];
    for (int d = 1; d < nd; ++d)
        if (sad[d] < minsad) { minsad = sad[d]; mind = d; }
    if (*s_tex < sp.texture) { *outp = FILTERED; return; }
    if (sp.uniq > 0) {
        int thresh = minsad + (minsad * sp.uniq / 100);
        for (int d = 0; d < nd; ++d)
            if ((d < mind - 1 || d > mind + 1) && sad[d] <= thresh) { *outp = FILTERED; return; }
    }
    sad[-1] = sad[1];
    sad[nd] = sad[nd - 2];
    int pp = sad[mind + 1], nn = sad[mind - 1];
    int den = pp + nn - 2 * sad[mind] + abs(pp - nn);
    int v = (nd - mind - 1) * 256 + (den != 0 ? (pp - nn) * 256 / den : 0) + 15;
    *outp = (short)(v >> 4);
}

// ---------------------------------------------------------------------------------------
// The same evaluation specialised for the reference's StereoBM::create(48, 21) (cap <= 63):
// raw patches arrive as aligned words, the 48 SADs run on packed bytes (vabsdiff4 with
// accumulate, as in stereo_dense.cu), winner / uniqueness are warp reductions instead of a
// serial tail. Same integers as the generic kernel above, ~3x fewer instructions.
constexpr int FS_ND = 48, FS_WIN = 21, FS_RAD = 10, FS_RW = FS_WIN + FS_ND - 1;   // 68
constexpr int FS_RAWH = FS_WIN + 2;                                              // 23
constexpr int FS_LP = 32, FS_RP = 80;        // byte pitch of the staged raw rows (8 / 20 words)
constexpr int FS_PLP = 24, FS_PRP = 72;      // byte pitch of the prefiltered rows (6 / 18 words)

__device__ __forceinline__ unsigned sad4_acc(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__global__ void __launch_bounds__(SB_THREADS)
stereo_bm_sparse_48_21_kernel(const uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride,
                              const uint8_t* __restrict__ right, size_t right_stride,
                              const float2* __restrict__ pts, int pts_stride,
                              short* __restrict__ disp16, SvoStereoParams sp,
                              const SvoSeqCall* __restrict__ call) {
    __shared__ __align__(16) unsigned s_rawL[FS_RAWH * FS_LP / 4];
    __shared__ __align__(16) unsigned s_rawR[FS_RAWH * FS_RP / 4];
    __shared__ __align__(16) unsigned s_PL[FS_WIN * FS_PLP / 4];
    __shared__ __align__(16) unsigned s_PR[FS_WIN * FS_PRP / 4];
    __shared__ int s_sadv[FS_ND + 2];
    __shared__ int s_texv;
    const SvoSeqCall cd = call[blockIdx.y];
    const int pi = blockIdx.x;
    if (pi >= cd.n_pts) return;
    const SvoLevel l0 = lay.lvl[0];
    const int W = l0.w, H = l0.h, pitch = l0.pitch;
    const uint8_t* __restrict__ L =
        pyr + (size_t)cd.seq * seq_stride + (size_t)cd.cur_slot * lay.slot_bytes + l0.off;
    const uint8_t* __restrict__ R = right + (size_t)cd.seq * right_stride;
    const float2 p = pts[(size_t)cd.seq * pts_stride + pi];
    short* outp = disp16 + (size_t)cd.seq * pts_stride + pi;
    const int x = (int)p.x, y = (int)p.y;     // float -> int truncation, image_processor.cpp:193
    const int cap = sp.cap;
    const short FILTERED = (short)(-16);
    const int tid = threadIdx.x;
    if (x < FS_ND - 1 + FS_RAD || x >= W - FS_RAD || y < FS_RAD || y >= H - FS_RAD) {
        if (tid == 0) *outp = FILTERED;
        return;
    }
    // ---- raw patches as aligned words. Column of raw index c: left xl0 + c, right xr0 + c
    // (xr0 may be -1: that column only feeds the prefilter of image column 0, which is `cap`
    // without looking; the staged words start at column max(xr0, 0) rounded down to 4)
    const int yl0 = y - FS_RAD - 1, xl0 = x - FS_RAD - 1, xr0 = x - FS_RAD - (FS_ND - 1) - 1;
    const int axl = xl0 & ~3, axr = max(xr0, 0) & ~3;
    const int offL = xl0 - axl, offR = xr0 - axr;           // offR may be -1
    for (int i = tid; i < FS_RAWH * (FS_LP / 4 + FS_RP / 4); i += SB_THREADS) {
        const int r = i / (FS_LP / 4 + FS_RP / 4), k = i - r * (FS_LP / 4 + FS_RP / 4);
        int gy = yl0 + r;
        gy = gy < 0 ? 1 : (gy > H - 1 ? H - 2 : gy);
        if (k < FS_LP / 4)
            s_rawL[r * (FS_LP / 4) + k] = __ldg(reinterpret_cast<const unsigned*>(L + (size_t)gy * pitch + axl) + k);
        else
            s_rawR[r * (FS_RP / 4) + (k - FS_LP / 4)] =
                __ldg(reinterpret_cast<const unsigned*>(R + (size_t)gy * pitch + axr) + (k - FS_LP / 4));
    }
    for (int i = tid; i < FS_WIN * FS_PLP / 4; i += SB_THREADS) s_PL[i] = 0u;
    for (int i = tid; i < FS_WIN * FS_PRP / 4; i += SB_THREADS) s_PR[i] = 0u;
    if (tid == 0) s_texv = 0;
    __syncthreads();
    // ---- XSOBEL prefilter (cap), bytes
    const unsigned char* rawL = reinterpret_cast<const unsigned char*>(s_rawL);
    const unsigned char* rawR = reinterpret_cast<const unsigned char*>(s_rawR);
    unsigned char* PL = reinterpret_cast<unsigned char*>(s_PL);
    unsigned char* PR = reinterpret_cast<unsigned char*>(s_PR);
    const bool odd_last = (H & 1);
    auto prefilter = [&](const unsigned char* raw, int rp, int off, int r, int c, int gx, int gy) -> int {
        if (gx == 0 || gx == W - 1 || (odd_last && gy == H - 1)) return cap;
        const unsigned char* a = raw + r * rp + off + c;     // row gy-1, col gx-1
        const unsigned char* b = a + rp;
        const unsigned char* d = b + rp;
        int v = ((int)a[2] - (int)a[0]) + 2 * ((int)b[2] - (int)b[0]) + ((int)d[2] - (int)d[0]);
        v = min(max(v, -cap), cap);
        return v + cap;
    };
    int tex = 0;
    for (int i = tid; i < FS_WIN * FS_WIN; i += SB_THREADS) {
        const int r = i / FS_WIN, c = i - r * FS_WIN;
        const int v = prefilter(rawL, FS_LP, offL, r, c, x - FS_RAD + c, y - FS_RAD + r);
        PL[r * FS_PLP + c] = (unsigned char)v;
        tex += abs(v - cap);
    }
    for (int i = tid; i < FS_WIN * FS_RW; i += SB_THREADS) {
        const int r = i / FS_RW, c = i - r * FS_RW;
        PR[r * FS_PRP + c] = (unsigned char)prefilter(rawR, FS_RP, offR, r, c,
                                                      x - FS_RAD - (FS_ND - 1) + c, y - FS_RAD + r);
    }
    tex = __reduce_add_sync(FULL, tex);
    if ((tid & 31) == 0) atomicAdd(&s_texv, tex);
    __syncthreads();
    // ---- SAD(D), D = tid < 48: 21 rows x (5 words + 1 byte), right operand at byte shift 47 - D
    if (tid < FS_ND) {
        const int shift = FS_ND - 1 - tid;
        const unsigned sh = (shift & 3) * 8;
        unsigned sad = 0;
#pragma unroll 3
        for (int r = 0; r < FS_WIN; ++r) {
            const unsigned* lw = s_PL + r * (FS_PLP / 4);
            const unsigned* rw = s_PR + r * (FS_PRP / 4) + (shift >> 2);
            unsigned w[7];
#pragma unroll
            for (int k = 0; k < 7; ++k) w[k] = (k + (shift >> 2) < FS_PRP / 4) ? rw[k] : 0u;
#pragma unroll
            for (int k = 0; k < 5; ++k) sad = sad4_acc(lw[k], __funnelshift_r(w[k], w[k + 1], sh), sad);
            sad = sad4_acc(lw[5] & 0xffu, __funnelshift_r(w[5], w[6], sh) & 0xffu, sad);
        }
        s_sadv[1 + shift] = (int)sad;          // internal order d = 47 - D
    }
    __syncthreads();
    // ---- winner (first minimum in internal order), texture, uniqueness, sub-pixel: warp 0
    if (tid >= 32) return;
    int* sad = s_sadv + 1;
    const int d0 = tid, d1 = tid + 32;
    const int v0 = sad[d0], v1 = d1 < FS_ND ? sad[d1] : 0x7fffffff;
    // key = sad * 64 + d: the minimum key is the first minimum (sad <= 441 * 126 < 2^16)
    unsigned key = min((unsigned)v0 * 64u + (unsigned)d0,
                       d1 < FS_ND ? (unsigned)v1 * 64u + (unsigned)d1 : 0xffffffffu);
    key = __reduce_min_sync(FULL, key);
    const int mind = (int)(key & 63u), minsad = (int)(key >> 6);
    bool reject = s_texv < sp.texture;
    if (sp.uniq > 0) {
        const int thresh = minsad + (minsad * sp.uniq / 100);
        const bool hit = ((d0 < mind - 1 || d0 > mind + 1) && v0 <= thresh) ||
                         (d1 < FS_ND && (d1 < mind - 1 || d1 > mind + 1) && v1 <= thresh);
        reject |= __any_sync(FULL, hit);
    }
    if (tid != 0) return;
    if (reject) { *outp = FILTERED; return; }
    sad[-1] = sad[1];
    sad[FS_ND] = sad[FS_ND - 2];
    const int pp = sad[mind + 1], nn = sad[mind - 1];
    const int den = pp + nn - 2 * sad[mind] + abs(pp - nn);
    const int v = (FS_ND - mind - 1) * 256 + (den != 0 ? (pp - nn) * 256 / den : 0) + 15;
    *outp = (short)(v >> 4);
}

// ---------------------------------------------------------------------------------------
constexpr int RP_THREADS = 256;

__device__ __forceinline__ int block_exclusive_scan_flag(bool flag, int* s_w, int& total) {
    // ordered compaction helper for RP_THREADS threads; returns this thread's output slot
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    unsigned bal = __ballot_sync(FULL, flag);
    if (lane == 0) s_w[wrp] = __popc(bal);
    __syncthreads();
    int off = 0;
    total = 0;
    for (int w = 0; w < RP_THREADS / 32; ++w) {
        int c = s_w[w];
        if (w < wrp) off += c;
        total += c;
    }
    __syncthreads();
    return off + __popc(bal & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(RP_THREADS)
reproject_kernel(const float2* __restrict__ pts, const short* __restrict__ disp16, int pts_stride,
                 float focal, float cx, float cy, float baseline, float2* __restrict__ valid2d,
                 float* __restrict__ pts3d, int* __restrict__ valid_idx, int* __restrict__ n_valid,
                 const SvoSeqCall* __restrict__ call) {
    __shared__ float s_m[16];
    __shared__ int s_w[RP_THREADS / 32];
    const SvoSeqCall cd = call[blockIdx.x];
    const size_t base = (size_t)cd.seq * pts_stride;
    if (threadIdx.x < 16) {
        // Q as the reference fills it (image_processor.cpp:183-189): doubles narrowed to f32
        float q[16];
        for (int i = 0; i < 16; ++i) q[i] = 0.f;
        q[0] = (float)(1.0 / (double)focal);
        q[5] = (float)(1.0 / (double)focal);
        q[3] = __fdiv_rn(-cx, focal);
        q[7] = __fdiv_rn(-cy, focal);
        q[11] = 1.f;
        q[14] = (float)(1.0 / (double)__fmul_rn(baseline, focal));
        int i = threadIdx.x >> 2, j = threadIdx.x & 3;
        double acc = 0.0;
        for (int k = 0; k < 4; ++k) acc = __dadd_rn(acc, __dmul_rn((double)cd.pose[i * 4 + k],
                                                                  (double)q[k * 4 + j]));
        s_m[threadIdx.x] = (float)acc;        // cv::gemm: double accumulate, f32 result
    }
    __syncthreads();
    int written = 0;
    for (int i0 = 0; i0 < cd.n_pts; i0 += RP_THREADS) {
        const int i = i0 + threadIdx.x;
        float2 p = make_float2(0.f, 0.f);
        float d = 0.f;
        if (i < cd.n_pts) {
            p = pts[base + i];
            d = __fmul_rn((float)disp16[base + i], 0.0625f);   // convertTo(CV_32F, 1/16)
        }
        const bool ok = (i < cd.n_pts) && (d > 0.f);
        int total;
        int slot = written + block_exclusive_scan_flag(ok, s_w, total);
        if (ok) {
            float v[4] = {p.x, p.y, d, 1.f};
            float w[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    acc = __dadd_rn(acc, __dmul_rn((double)s_m[r * 4 + k], (double)v[k]));
                w[r] = (float)acc;
            }
            valid2d[base + slot] = p;
            valid_idx[base + slot] = i;
            pts3d[(base + slot) * 3 + 0] = __fdiv_rn(w[0], w[3]);
            pts3d[(base + slot) * 3 + 1] = __fdiv_rn(w[1], w[3]);
            pts3d[(base + slot) * 3 + 2] = __fdiv_rn(w[2], w[3]);
        }
        written += total;
    }
    if (threadIdx.x == 0) n_valid[cd.seq] = written;
}

// dist < md with dist = sqrt_rn(d2) is a threshold test on d2 itself (sqrt_rn is monotonic):
// the smallest float t with sqrt_rn(t) >= md. Found by walking a few ulps around md*md.
__device__ __forceinline__ float dist2_threshold(float md) {
    if (!(md > 0.f)) return 0.f;                      // dist >= 0: nothing is closer than md
    float t = __fmul_rn(md, md);
    for (int k = 0; k < 8 && t > 0.f; ++k) {
        const float lo = __int_as_float(__float_as_int(t) - 1);
        if (__fsqrt_rn(lo) >= md) t = lo; else break;
    }
    for (int k = 0; k < 8 && __fsqrt_rn(t) < md; ++k) t = __int_as_float(__float_as_int(t) + 1);
    return t;
}

constexpr int DD_BLOCKS = 4;      // blocks per sequence
constexpr int DD_TILE = 1024;     // tracked points staged per pass

// New detections closer than min_distance to a tracked inlier are dropped, order kept
// (src/image_processor.cpp:118-131). DD_BLOCKS blocks of a sequence test a quarter of the
// detections each against the tracked points (staged in shared memory) and leave a flag per
// detection in keep_idx; the block that finishes last compacts. `ticket` is zero between calls.
__global__ void __launch_bounds__(RP_THREADS)
dedupe_kernel(const float2* __restrict__ det, const int* __restrict__ n_det, int det_stride,
              const float2* __restrict__ inl, int inl_stride, int inl_buffered,
              float min_distance, float2* __restrict__ out, int* keep_idx,
              int* __restrict__ n_out, int out_stride, SvoSeqCall* call, unsigned* ticket) {
    __shared__ float2 s_q[DD_TILE];
    __shared__ int s_w[RP_THREADS / 32];
    __shared__ bool s_last;
    const SvoSeqCall cd = call[blockIdx.x];
    if (cd.flags & SVO_CALL_SKIP) return;
    const float2* d = det + (size_t)cd.seq * det_stride;
    // inl_buffered > 0: the tracked set is the tracker's own double-buffered feature array
    const float2* q = inl + (size_t)cd.seq * inl_stride + (size_t)cd.feat_buf * inl_buffered;
    const int nd = n_det ? n_det[cd.seq] : cd.n_pts;
    const int ni = cd.n_aux;
    const size_t ob = (size_t)cd.seq * out_stride;
    const float thr = dist2_threshold(min_distance);
    const int per = (nd + DD_BLOCKS - 1) / DD_BLOCKS;
    const int i_beg = blockIdx.y * per, i_end = min(nd, i_beg + per);
    for (int i0 = i_beg; i0 < i_end; i0 += RP_THREADS) {
        const int i = i0 + threadIdx.x;
        const float2 p = (i < i_end) ? d[i] : make_float2(0.f, 0.f);
        bool near = false;
        for (int j0 = 0; j0 < ni; j0 += DD_TILE) {
            const int nj = min(DD_TILE, ni - j0);
            __syncthreads();
            for (int j = threadIdx.x; j < nj; j += RP_THREADS) s_q[j] = q[j0 + j];
            __syncthreads();
#pragma unroll 8
            for (int j = 0; j < nj; ++j) {
                const float2 t = s_q[j];
                const float dx = __fsub_rn(p.x, t.x), dy = __fsub_rn(p.y, t.y);
                near |= __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < thr;
            }
        }
        if (i < i_end) keep_idx[ob + i] = near ? 0 : 1;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(ticket + cd.seq, 1u) == DD_BLOCKS - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    int written = 0;
    for (int i0 = 0; i0 < nd; i0 += RP_THREADS) {
        // flags of this chunk are read before any slot (<= i) of it is written: in place is safe
        const int i = i0 + threadIdx.x;
        const bool keep = i < nd && __ldcg(keep_idx + ob + i) != 0;
        const float2 p = keep ? d[i] : make_float2(0.f, 0.f);
        int total;
        int slot = written + block_exclusive_scan_flag(keep, s_w, total);
        if (keep) {
            out[ob + slot] = p;
            keep_idx[ob + slot] = i;
        }
        written += total;
    }
    if (threadIdx.x == 0) {
        n_out[cd.seq] = written;
        call[blockIdx.x].n_pts = written;      // the stereo stage of the same call reads it
        ticket[cd.seq] = 0;
    }
}

}  // namespace

static size_t stereo_smem_bytes(const SvoStereoParams& sp) {
    int win = sp.win, nd = sp.ndisp, rw = win + nd - 1;
    size_t b = (size_t)(win + 2) * (win + 2) + (size_t)(win + 2) * (rw + 2) + (size_t)win * win +
               (size_t)win * rw;
    b = (b + 15) & ~(size_t)15;
    return b + (size_t)(nd + 2 + 1) * sizeof(int);
}

cudaError_t svo_launch_stereo_bm_sparse(const uint8_t* pyr, const SvoPyrLayout& lay,
                                        size_t seq_stride, const uint8_t* right,
                                        size_t right_stride, const float2* pts, int pts_stride,
                                        short* disp16, SvoStereoParams sp, int max_pts,
                                        const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    if (max_pts <= 0) return cudaSuccess;
    static const bool generic_only = getenv("SVO_STEREO_GENERIC") && atoi(getenv("SVO_STEREO_GENERIC")) != 0;
    if (sp.ndisp == FS_ND && sp.win == FS_WIN && sp.cap >= 1 && sp.cap <= 63 && !generic_only)
        stereo_bm_sparse_48_21_kernel<<<dim3(max_pts, n_call), SB_THREADS, 0, st>>>(
            pyr, lay, seq_stride, right, right_stride, pts, pts_stride, disp16, sp, d_call);
    else
        stereo_bm_sparse_kernel<<<dim3(max_pts, n_call), SB_THREADS, stereo_smem_bytes(sp), st>>>(
            pyr, lay, seq_stride, right, right_stride, pts, pts_stride, disp16, sp, d_call);
    return cudaGetLastError();
}

cudaError_t svo_launch_reproject(const float2* pts, const short* disp16, int pts_stride,
                                 float focal, float cx, float cy, float baseline, float2* valid2d,
                                 float* pts3d, int* valid_idx, int* n_valid,
                                 const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    reproject_kernel<<<n_call, RP_THREADS, 0, st>>>(pts, disp16, pts_stride, focal, cx, cy,
                                                    baseline, valid2d, pts3d, valid_idx, n_valid,
                                                    d_call);
    return cudaGetLastError();
}

cudaError_t svo_launch_dedupe(const float2* det, const int* n_det, int det_stride,
                              const float2* inl, int inl_stride, int inl_buffered,
                              float min_distance, float2* out,
                              int* keep_idx, int* n_out, int out_stride,
                              SvoSeqCall* d_call, unsigned* ticket, int n_call, cudaStream_t st) {
    dedupe_kernel<<<dim3(n_call, DD_BLOCKS), RP_THREADS, 0, st>>>(
        det, n_det, det_stride, inl, inl_stride, inl_buffered, min_distance, out, keep_idx, n_out,
        out_stride, d_call, ticket);
    return cudaGetLastError();
}
