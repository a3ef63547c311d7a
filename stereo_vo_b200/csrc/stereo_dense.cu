// stereo_bm_dense: the full disparity map of cv::StereoBM::create(48, 21)->compute(left, right)
// (reference call site src/image_processor.cpp:173-176). Same arithmetic
// as stereo_bm_sparse (stereo.cu; pinned in oracle/restate.py:stereo_bm_at, Appendix A2), for
// every pixel: XSOBEL prefilter (cap 31), 48 SADs over 21x21, first minimum in internal order
// d = 47 - D, texture threshold, uniqueness ratio, parabola sub-pixel, CV_16S in 1/16 px;
// pixels outside x in [57, W-10), y in [10, H-10) hold (minDisparity - 1) << 4 = -16.
//
// One kernel, raw u8 images in, int16 map out (algorithmic bytes 4 N: N + N read, 2 N written):
//   * a block owns 32 output columns x a band of <= DN_BAND output rows. It prefilters the
//     (band + 20) x (32 + 20 [+ 47 for the right image]) pixels it needs straight from the raw
//     images into shared memory, once;
//   * thread (lane = column, warp k = disparity class) carries SAD(x, y, D) for the twelve
//     D = k + 4 j in registers and slides the 21-row window down the band: the entering row's
//     21-column sums are added, the leaving row's subtracted. D in steps of 4 makes the right
//     image operand of disparity D + 4 the SAME aligned words one word further left, so a
//     row costs 25 shared-memory words + 23 funnel shifts, then 6 VABSDIFF4.ACC (4 absolute
//     differences + accumulate per instruction) + 1 mask per disparity;
//   * the window's texture sum is the 49th "disparity" (left operand against the cap bytes);
//   * every 4 output rows the SADs of the 4 warps meet in shared memory and thread
//     (column, row) runs the winner / uniqueness / sub-pixel tail of
//     findStereoCorrespondenceBM.
// Integer throughout, bit-exact. Bound: SM issue (about 1 100 lane-instructions per pixel);
// DRAM traffic is the 4 N of the interface (tiles overlap only through L2).
#include "svo_common.cuh"

namespace {

constexpr int DN_ND = 48;          // numDisparities of the reference (image_processor.cpp:174)
constexpr int DN_WIN = 21;         // SAD window
constexpr int DN_RAD = DN_WIN / 2;
constexpr int DN_TW = 32;          // output columns per block
constexpr int DN_BAND = 64;        // output rows per block
constexpr int DN_PLW = 56;         // prefiltered left tile row: 32 + 20 columns, + a spare word
constexpr int DN_PRW = 104;        // prefiltered right tile row: 32 + 20 + 47 columns, padded
constexpr int DN_ROWS = DN_BAND + DN_WIN - 1;
constexpr int DN_NJ = DN_ND / 4;   // disparities per thread
constexpr int DN_SADW = DN_ND + 2; // SAD list with one pad entry either side
constexpr short DN_FILTERED = (short)(-16);

// prefilterXSobel at (gx, gy): borders as OpenCV walks them (same rules as stereo.cu)
__device__ __forceinline__ int xsobel(const uint8_t* __restrict__ img, int pitch, int W, int H,
                                      int gx, int gy, int cap) {
    if (gx <= 0 || gx >= W - 1 || ((H & 1) && gy == H - 1)) return cap;
    const int ya = gy - 1 < 0 ? 1 : gy - 1;
    const int yc = gy + 1 > H - 1 ? H - 2 : gy + 1;
    const uint8_t* a = img + (size_t)ya * pitch + gx;
    const uint8_t* b = img + (size_t)gy * pitch + gx;
    const uint8_t* c = img + (size_t)yc * pitch + gx;
    int v = ((int)__ldg(a + 1) - (int)__ldg(a - 1)) + 2 * ((int)__ldg(b + 1) - (int)__ldg(b - 1)) +
            ((int)__ldg(c + 1) - (int)__ldg(c - 1));
    v = min(max(v, -cap), cap);
    return v + cap;
}

struct RowOps {
    unsigned l[6];                 // 21 left bytes (l[5]: byte 20 only)
    unsigned r[17];                // 65 + 3 right bytes starting at column x - 10 - (k + 44)
};

__device__ __forceinline__ void load_row(RowOps& o, const unsigned* __restrict__ pl,
                                         const unsigned* __restrict__ pr, int lw, int ls, int rw,
                                         int rs) {
    unsigned w[18];
#pragma unroll
    for (int i = 0; i < 7; ++i) w[i] = pl[lw + i];
#pragma unroll
    for (int i = 0; i < 6; ++i) o.l[i] = __funnelshift_r(w[i], w[i + 1], ls);
    o.l[5] &= 0xffu;
#pragma unroll
    for (int i = 0; i < 18; ++i) w[i] = pr[rw + i];
#pragma unroll
    for (int i = 0; i < 17; ++i) o.r[i] = __funnelshift_r(w[i], w[i + 1], rs);
}

template <int SIGN>
__device__ __forceinline__ void accumulate(int (&sad)[DN_NJ], int& tex, const RowOps& o,
                                           unsigned capw) {
#pragma unroll
    for (int j = 0; j < DN_NJ; ++j) {
        // D = k + 4 j: the operand starts 4 (11 - j) bytes into the stream
        unsigned s = 0;
#pragma unroll
        for (int i = 0; i < 5; ++i) s = __vsadu4(o.l[i], o.r[11 - j + i]) + s;
        s = __vsadu4(o.l[5], o.r[16 - j] & 0xffu) + s;
        sad[j] += SIGN * (int)s;
    }
    unsigned t = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) t = __vsadu4(o.l[i], capw) + t;
    t = __vsadu4(o.l[5], capw & 0xffu) + t;
    tex += SIGN * (int)t;
}

__global__ void __launch_bounds__(128)
stereo_bm_dense_kernel(const uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride,
                       const uint8_t* __restrict__ right, size_t right_stride,
                       short* __restrict__ disp, int dpitch, size_t disp_stride, int cap,
                       int texture, int uniq, const SvoSeqCall* __restrict__ call) {
    __shared__ __align__(16) unsigned char s_pl[DN_ROWS * DN_PLW];
    __shared__ __align__(16) unsigned char s_pr[DN_ROWS * DN_PRW];
    __shared__ unsigned short s_sad[4][DN_SADW][DN_TW];
    __shared__ int s_tex[4][DN_TW];

    const SvoSeqCall cd = call[blockIdx.z];
    const SvoLevel l0 = lay.lvl[0];
    const int W = l0.w, H = l0.h, pitch = l0.pitch;
    const uint8_t* __restrict__ L =
        pyr + (size_t)cd.seq * seq_stride + (size_t)cd.cur_slot * lay.slot_bytes + l0.off;
    const uint8_t* __restrict__ R = right + (size_t)cd.seq * right_stride;
    short* __restrict__ out = disp + (size_t)cd.seq * disp_stride;

    const int xt = DN_ND - 1 + DN_RAD + blockIdx.x * DN_TW;      // first output column
    const int yt = DN_RAD + blockIdx.y * DN_BAND;                // first output row
    const int nrows = min(DN_BAND, H - DN_RAD - yt);             // output rows of this band
    const int tid = threadIdx.x, lane = tid & 31, k = tid >> 5;

    // ---- prefiltered tiles: rows yt - 10 .. yt + nrows + 9 (all inside the image)
    const int trows = nrows + DN_WIN - 1;
    for (int i = tid; i < trows * DN_PLW; i += 128) {
        const int r = i / DN_PLW, c = i - r * DN_PLW;
        const int gx = min(xt - DN_RAD + c, W - 1);              // past W-1: never reaches a valid pixel
        s_pl[i] = (unsigned char)xsobel(L, pitch, W, H, gx, yt - DN_RAD + r, cap);
    }
    for (int i = tid; i < trows * DN_PRW; i += 128) {
        const int r = i / DN_PRW, c = i - r * DN_PRW;
        const int gx = min(xt - DN_RAD - (DN_ND - 1) + c, W - 1);
        s_pr[i] = (unsigned char)xsobel(R, pitch, W, H, gx, yt - DN_RAD + r, cap);
    }
    __syncthreads();

    // per-thread operand addressing: left bytes start at tile column `lane`, the right stream
    // at tile column lane + 3 - k (image column x - 10 - (k + 44))
    const int lw = lane >> 2, ls = (lane & 3) * 8;
    const int rb = lane + 3 - k;
    const int rw = rb >> 2, rs = (rb & 3) * 8;
    const unsigned capw = (unsigned)cap * 0x01010101u;
    const unsigned* pl32 = reinterpret_cast<const unsigned*>(s_pl);
    const unsigned* pr32 = reinterpret_cast<const unsigned*>(s_pr);

    int sad[DN_NJ];
#pragma unroll
    for (int j = 0; j < DN_NJ; ++j) sad[j] = 0;
    int tex = 0;
    RowOps o;
    for (int r = 0; r < DN_WIN - 1; ++r) {
        load_row(o, pl32 + r * (DN_PLW / 4), pr32 + r * (DN_PRW / 4), lw, ls, rw, rs);
        accumulate<1>(sad, tex, o, capw);
    }
    for (int y0 = 0; y0 < nrows; y0 += 4) {
        const int ng = min(4, nrows - y0);
        for (int g = 0; g < ng; ++g) {
            const int ry = y0 + g;                               // output row yt + ry
            load_row(o, pl32 + (ry + DN_WIN - 1) * (DN_PLW / 4),
                     pr32 + (ry + DN_WIN - 1) * (DN_PRW / 4), lw, ls, rw, rs);
            accumulate<1>(sad, tex, o, capw);
#pragma unroll
            for (int j = 0; j < DN_NJ; ++j)                      // internal order d = 47 - D
                s_sad[g][1 + (DN_ND - 1 - (k + 4 * j))][lane] = (unsigned short)sad[j];
            if (k == 0) s_tex[g][lane] = tex;
            load_row(o, pl32 + ry * (DN_PLW / 4), pr32 + ry * (DN_PRW / 4), lw, ls, rw, rs);
            accumulate<-1>(sad, tex, o, capw);
        }
        __syncthreads();
        // ---- tail of findStereoCorrespondenceBM: thread (lane, k) finishes row y0 + k
        const int x = xt + lane, y = yt + y0 + k;
        if (k < ng && x < W - DN_RAD) {
            short res;
            int mind = 0, minsad = s_sad[k][1][lane];
#pragma unroll 8
            for (int d = 1; d < DN_ND; ++d) {
                const int v = s_sad[k][1 + d][lane];
                if (v < minsad) { minsad = v; mind = d; }
            }
            bool ok = s_tex[k][lane] >= texture;
            if (ok && uniq > 0) {
                const int thresh = minsad + (minsad * uniq / 100);
#pragma unroll 8
                for (int d = 0; d < DN_ND; ++d) {
                    const int v = s_sad[k][1 + d][lane];
                    if ((d < mind - 1 || d > mind + 1) && v <= thresh) ok = false;
                }
            }
            if (ok) {
                const int c0 = s_sad[k][1 + mind][lane];
                const int pp = (mind + 1 < DN_ND) ? s_sad[k][2 + mind][lane] : s_sad[k][DN_ND - 1][lane];
                const int nn = (mind > 0) ? s_sad[k][mind][lane] : s_sad[k][2][lane];
                const int den = pp + nn - 2 * c0 + abs(pp - nn);
                const int v = (DN_ND - mind - 1) * 256 + (den != 0 ? (pp - nn) * 256 / den : 0) + 15;
                res = (short)(v >> 4);
            } else {
                res = DN_FILTERED;
            }
            out[(size_t)y * dpitch + x] = res;
        }
        __syncthreads();
    }
}

// the frame of pixels StereoBM never evaluates
__global__ void stereo_bm_dense_border_kernel(short* __restrict__ disp, int dpitch,
                                              size_t disp_stride, int W, int H,
                                              const SvoSeqCall* __restrict__ call) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W) return;
    const bool inside = x >= DN_ND - 1 + DN_RAD && x < W - DN_RAD && y >= DN_RAD && y < H - DN_RAD;
    if (!inside) disp[(size_t)call[blockIdx.z].seq * disp_stride + (size_t)y * dpitch + x] = DN_FILTERED;
}

}  // namespace

bool svo_stereo_dense_supported(const SvoStereoParams& sp) {
    return sp.ndisp == DN_ND && sp.win == DN_WIN && sp.cap >= 1 && sp.cap <= 63;
}

cudaError_t svo_launch_stereo_bm_dense(const uint8_t* pyr, const SvoPyrLayout& lay,
                                       size_t seq_stride, const uint8_t* right,
                                       size_t right_stride, short* disp, int dpitch,
                                       size_t disp_stride, SvoStereoParams sp,
                                       const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    const int W = lay.lvl[0].w, H = lay.lvl[0].h;
    stereo_bm_dense_border_kernel<<<dim3((W + 255) / 256, H, n_call), 256, 0, st>>>(
        disp, dpitch, disp_stride, W, H, d_call);
    const int vw = W - DN_RAD - (DN_ND - 1 + DN_RAD), vh = H - 2 * DN_RAD;
    if (vw > 0 && vh > 0) {
        dim3 grid((vw + DN_TW - 1) / DN_TW, (vh + DN_BAND - 1) / DN_BAND, n_call);
        stereo_bm_dense_kernel<<<grid, 128, 0, st>>>(pyr, lay, seq_stride, right, right_stride,
                                                     disp, dpitch, disp_stride, sp.cap, sp.texture,
                                                     sp.uniq, d_call);
    }
    return cudaGetLastError();
}
