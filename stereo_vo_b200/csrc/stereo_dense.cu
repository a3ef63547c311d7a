// stereo_bm_dense: the full disparity map of cv::StereoBM::create(48, 21)->compute(left, right)
// (reference call site src/image_processor.cpp:173-176). Same arithmetic
// as stereo_bm_sparse (stereo.cu; pinned in oracle/restate.py:stereo_bm_at, Appendix A2), for
// every pixel: XSOBEL prefilter (cap 31), 48 SADs over 21x21, first minimum in internal order
// d = 47 - D, texture threshold, uniqueness ratio, parabola sub-pixel, CV_16S in 1/16 px;
// pixels outside x in [57, W-10), y in [10, H-10) hold (minDisparity - 1) << 4 = -16.
//
// One kernel, raw u8 images in, int16 map out (algorithmic bytes 4 N: N + N read, 2 N written):
//   * a block owns 32 output columns x a band of output rows (DN_BAND for a lone image, equal
//     bands of <= DN_BAND_TALL when the launch holds many). It prefilters the
//     (band + 20) x (32 + 20 [+ 47 for the right image]) pixels it needs straight from the raw
//     images into shared memory, once;
//   * thread (lane = column, warp k = disparity class) carries SAD(x, y, D) for the twelve
//     D = k + 4 j in registers, two per word in 16-bit halves, and slides the 21-row window
//     down the band: the entering row's 21-column sums are added and parked in a per-thread
//     ring of 21 rows in shared memory, the leaving row's are read back from it and
//     subtracted (nothing is computed twice). D in steps of 4 makes the right image operand
//     of disparity D + 4 the SAME aligned words one word further left, so a row costs 25
//     shared-memory words + 23 funnel shifts, then 6 VABSDIFF4.ACC (4 absolute differences +
//     accumulate per instruction) + 1 mask per disparity, 6 adds, 6 subtracts;
//   * the window's texture sum is kept the same way by warp 0 (left operand against the cap);
//   * every 4 output rows the SADs of the 4 warps meet in shared memory and thread
//     (column, row) runs the winner / uniqueness / sub-pixel tail of
//     findStereoCorrespondenceBM.
// Integer throughout, bit-exact. Bound: the integer ALU pipe (about 500 lane-instructions per
// pixel); DRAM traffic is the 4 N of the interface (tiles overlap only through L2).
#include "svo_common.cuh"

namespace {

constexpr int DN_ND = 48;          // numDisparities of the reference (image_processor.cpp:174)
constexpr int DN_WIN = 21;         // SAD window
constexpr int DN_RAD = DN_WIN / 2;
constexpr int DN_TW = 32;          // output columns per block
constexpr int DN_BAND = 64;        // output rows per block (few images in the launch)
constexpr int DN_BAND_TALL = 184;  // ... upper bound when the launch holds many images (2 blocks per SM)
constexpr int DN_PLW = 56;         // prefiltered left tile row: 32 + 20 columns, + a spare word
constexpr int DN_PRW = 104;        // prefiltered right tile row: 32 + 20 + 47 columns, padded
constexpr int DN_NJ = DN_ND / 4;   // disparities per thread
constexpr short DN_FILTERED = (short)(-16);

// prefilterXSobel for a 4 x 4 patch: columns gx0 .. gx0 + 3 (one word per row) of rows
// gy0 .. gy0 + 3, from the 6 x 6 raw pixels around it (36 independent byte loads in flight).
// Borders as OpenCV walks them (same rules as stereo.cu): columns 0 and W-1 and the last row of
// an odd-height image hold the cap; rows above / below the image reflect onto rows 1 / H-2.
// Columns past W-1 are clamped for the loads and forced to the cap: they never reach a valid
// output pixel. Rows past H-1 (only in the last, partial patch of a tile) are computed from
// clamped rows and dropped by the caller.
__device__ __forceinline__ void xsobel4x4(unsigned (&out)[4], const uint8_t* __restrict__ img,
                                          int pitch, int W, int H, int gx0, int gy0, int cap) {
    int raw[6][6];
    int col[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) col[i] = min(max(gx0 - 1 + i, 0), W - 1);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
        int y = gy0 - 1 + q;                        // raw row; tile rows themselves are inside the image
        y = y < 0 ? 1 : (y == H ? H - 2 : (y > H ? H - 1 : y));   // rows -1 and H reflect; later ones unused
        const uint8_t* rowp = img + (size_t)y * pitch;
#pragma unroll
        for (int i = 0; i < 6; ++i) raw[q][i] = (int)__ldg(rowp + col[i]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int gy = gy0 + r;
        const bool row_cap = (H & 1) && gy == H - 1;
        int s[6];                                   // vertical 1-2-1 sums of columns gx0-1 .. gx0+4
#pragma unroll
        for (int i = 0; i < 6; ++i) s[i] = raw[r][i] + 2 * raw[r + 1][i] + raw[r + 2][i];
        unsigned w = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int gx = gx0 + t;
            int v = min(max(s[t + 2] - s[t], -cap), cap) + cap;
            if (row_cap || gx <= 0 || gx >= W - 1) v = cap;
            w |= (unsigned)v << (8 * t);
        }
        out[r] = w;
    }
}

struct RowOps {
    unsigned l[6];                 // 21 left bytes (l[5]: byte 20 only)
    unsigned r[17];                // 65 + 3 right bytes starting at column x - 10 - (k + 44)
};

struct RawRow {
    unsigned wl[7], wr[18];        // the aligned words that hold the operands
};

__device__ __forceinline__ void load_raw(RawRow& w, const unsigned* __restrict__ pl,
                                         const unsigned* __restrict__ pr, int lw, int rw) {
#pragma unroll
    for (int i = 0; i < 7; ++i) w.wl[i] = pl[lw + i];
#pragma unroll
    for (int i = 0; i < 18; ++i) w.wr[i] = pr[rw + i];
}

__device__ __forceinline__ void make_ops(RowOps& o, const RawRow& w, int ls, int rs) {
#pragma unroll
    for (int i = 0; i < 6; ++i) o.l[i] = __funnelshift_r(w.wl[i], w.wl[i + 1], ls);
    o.l[5] &= 0xffu;
#pragma unroll
    for (int i = 0; i < 17; ++i) o.r[i] = __funnelshift_r(w.wr[i], w.wr[i + 1], rs);
}

// |a - b| summed over the four bytes, + c: one VABSDIFF4.U8.ACC (written as PTX so that the
// accumulator operand is used; `__vsadu4(a, b) + c` compiles to a zero-accumulator op and adds)
__device__ __forceinline__ unsigned sad4(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// 21-column sums of one row for the thread's twelve disparities, two per word (16-bit halves:
// a row sum is <= 21 * 126, a window sum <= 441 * 126 < 2^16, and sums of halves never carry)
__device__ __forceinline__ void row_sums(unsigned (&hp)[DN_NJ / 2], const RowOps& o) {
#pragma unroll
    for (int jj = 0; jj < DN_NJ / 2; ++jj) {
        // D = k + 4 j: the operand starts 4 (11 - j) bytes into the stream
        const int j0 = 2 * jj, j1 = 2 * jj + 1;
        unsigned a = 0, b = 0;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            a = sad4(o.l[i], o.r[11 - j0 + i], a);
            b = sad4(o.l[i], o.r[11 - j1 + i], b);
        }
        a = sad4(o.l[5], o.r[16 - j0] & 0xffu, a);
        b = sad4(o.l[5], o.r[16 - j1] & 0xffu, b);
        hp[jj] = a + b * 65536u;
    }
}

__device__ __forceinline__ unsigned row_texture(const RowOps& o, unsigned capw) {
    unsigned t = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) t = sad4(o.l[i], capw, t);
    return sad4(o.l[5], capw & 0xffu, t);
}

// shared-memory carve-up (dynamic): ring | sad | texture ring | texture | left tile | right tile
constexpr int DN_RING_WORDS = DN_WIN * (DN_NJ / 2) * 128;      // per-thread row sums of 21 rows
constexpr int DN_SAD_WORDS = 4 * 4 * (DN_NJ / 2) * DN_TW;      // [row of group][k][jj][lane]
constexpr int DN_TEXR_WORDS = DN_WIN * DN_TW;
constexpr int DN_TEX_WORDS = 4 * DN_TW;

__global__ void __launch_bounds__(128, 2)
stereo_bm_dense_kernel(const uint8_t* __restrict__ pyr, SvoPyrLayout lay, size_t seq_stride,
                       const uint8_t* __restrict__ right, size_t right_stride,
                       short* __restrict__ disp, int dpitch, size_t disp_stride, int band, int cap,
                       int texture, int uniq, const SvoSeqCall* __restrict__ call) {
    extern __shared__ __align__(16) unsigned s_dyn[];
    unsigned* s_ring = s_dyn;
    unsigned* s_sad = s_ring + DN_RING_WORDS;
    unsigned* s_texr = s_sad + DN_SAD_WORDS;
    int* s_tex = reinterpret_cast<int*>(s_texr + DN_TEXR_WORDS);
    unsigned* s_pl = reinterpret_cast<unsigned*>(s_tex + DN_TEX_WORDS);
    unsigned* s_pr = s_pl + (band + DN_WIN - 1) * (DN_PLW / 4);

    const SvoSeqCall cd = call[blockIdx.z];
    const SvoLevel l0 = lay.lvl[0];
    const int W = l0.w, H = l0.h, pitch = l0.pitch;
    const uint8_t* __restrict__ L =
        pyr + (size_t)cd.seq * seq_stride + (size_t)cd.cur_slot * lay.slot_bytes + l0.off;
    const uint8_t* __restrict__ R = right + (size_t)cd.seq * right_stride;
    short* __restrict__ out = disp + (size_t)cd.seq * disp_stride;

    const int xt = DN_ND - 1 + DN_RAD + blockIdx.x * DN_TW;      // first output column
    const int yt = DN_RAD + blockIdx.y * band;                   // first output row
    const int nrows = min(band, H - DN_RAD - yt);                // output rows of this band
    const int tid = threadIdx.x, lane = tid & 31, k = tid >> 5;

    // ---- prefiltered tiles: rows yt - 10 .. yt + nrows + 9 (all inside the image)
    const int trows = nrows + DN_WIN - 1;
    const int tquads = (trows + 3) >> 2;
    for (int i = tid; i < tquads * (DN_PLW / 4 + DN_PRW / 4); i += 128) {
        const int q = i / (DN_PLW / 4 + DN_PRW / 4), c = i - q * (DN_PLW / 4 + DN_PRW / 4);
        const bool left = c < DN_PLW / 4;
        const int cw = left ? c : c - DN_PLW / 4;
        unsigned w4[4];
        xsobel4x4(w4, left ? L : R, pitch, W, H,
                  (left ? xt - DN_RAD : xt - DN_RAD - (DN_ND - 1)) + 4 * cw, yt - DN_RAD + 4 * q, cap);
        unsigned* dst = left ? s_pl + cw : s_pr + cw;
        const int rwords = left ? DN_PLW / 4 : DN_PRW / 4;
#pragma unroll
        for (int r = 0; r < 4; ++r)
            if (4 * q + r < trows) dst[(4 * q + r) * rwords] = w4[r];
    }
    __syncthreads();

    // per-thread operand addressing: left bytes start at tile column `lane`, the right stream
    // at tile column lane + 3 - k (image column x - 10 - (k + 44))
    const int lw = lane >> 2, ls = (lane & 3) * 8;
    const int rb = lane + 3 - k;
    const int rw = rb >> 2, rs = (rb & 3) * 8;
    const unsigned capw = (unsigned)cap * 0x01010101u;
    unsigned* ring = s_ring + tid;                               // [slot][jj] 128 words apart

    unsigned sadp[DN_NJ / 2], hp[DN_NJ / 2];
#pragma unroll
    for (int jj = 0; jj < DN_NJ / 2; ++jj) sadp[jj] = 0;
    unsigned tex = 0;
    RowOps o;
    RawRow cur, nxt;               // the next row's words are fetched while this row is summed
    const unsigned* pl = s_pl + lw;
    const unsigned* pr = s_pr + rw;
    load_raw(cur, pl, pr, 0, 0);
#pragma unroll 2
    for (int r = 0; r < DN_WIN - 1; ++r) {                       // window rows 0 .. 19 -> slots 0 .. 19
        load_raw(nxt, pl + (r + 1) * (DN_PLW / 4), pr + (r + 1) * (DN_PRW / 4), 0, 0);
        make_ops(o, cur, ls, rs);
        row_sums(hp, o);
#pragma unroll
        for (int jj = 0; jj < DN_NJ / 2; ++jj) {
            sadp[jj] += hp[jj];
            ring[(r * (DN_NJ / 2) + jj) * 128] = hp[jj];
        }
        if (k == 0) {
            const unsigned t = row_texture(o, capw);
            tex += t;
            s_texr[r * DN_TW + lane] = t;
        }
        cur = nxt;
    }
    int slot_in = DN_WIN - 1, slot_out = 0;                      // ring slots of rows ry + 20, ry
    for (int y0 = 0; y0 < nrows; y0 += 4) {
        const int ng = min(4, nrows - y0);
#pragma unroll
        for (int g = 0; g < 4; ++g) {                            // unrolled: cur / nxt swap by renaming
            if (g >= ng) break;
            const int ry = y0 + g;                               // output row yt + ry; cur = tile row ry + 20
            if (ry + 1 < nrows)
                load_raw(nxt, pl + (ry + DN_WIN) * (DN_PLW / 4), pr + (ry + DN_WIN) * (DN_PRW / 4), 0, 0);
            unsigned old[DN_NJ / 2], told = 0;                   // row ry leaves: slot_out != slot_in
#pragma unroll
            for (int jj = 0; jj < DN_NJ / 2; ++jj) old[jj] = ring[(slot_out * (DN_NJ / 2) + jj) * 128];
            if (k == 0) told = s_texr[slot_out * DN_TW + lane];
            make_ops(o, cur, ls, rs);
            row_sums(hp, o);
#pragma unroll
            for (int jj = 0; jj < DN_NJ / 2; ++jj) {
                sadp[jj] += hp[jj];                              // window complete: rows ry .. ry + 20
                ring[(slot_in * (DN_NJ / 2) + jj) * 128] = hp[jj];
                s_sad[((g * 4 + k) * (DN_NJ / 2) + jj) * DN_TW + lane] = sadp[jj];
                sadp[jj] -= old[jj];
            }
            if (k == 0) {
                const unsigned t = row_texture(o, capw);
                tex += t;
                s_texr[slot_in * DN_TW + lane] = t;
                s_tex[g * DN_TW + lane] = (int)tex;
                tex -= told;
            }
            slot_in = (slot_in == DN_WIN - 1) ? 0 : slot_in + 1;
            slot_out = (slot_out == DN_WIN - 1) ? 0 : slot_out + 1;
            cur = nxt;
        }
        __syncthreads();
        // ---- tail of findStereoCorrespondenceBM: thread (lane, k) finishes row y0 + k.
        // SAD of internal disparity d (= 47 - D, D = kk + 4 j) sits in half j & 1 of word
        // [row][kk][j >> 1][lane]
        const int x = xt + lane, y = yt + y0 + k;
        if (k < ng && x < W - DN_RAD) {
            // this thread's 48 SADs: word (kk, jj) of row k, 32 words apart; nobody else reads
            // them, and the next group's sums arrive only after the barrier below
            unsigned* sw = s_sad + (k * 4) * (DN_NJ / 2) * DN_TW + lane;
            auto word_of = [&](int d) -> int {
                const int D = DN_ND - 1 - d;
                return ((D & 3) * (DN_NJ / 2) + (D >> 3)) * DN_TW;
            };
            auto sad_at = [&](int d) -> int {
                const int D = DN_ND - 1 - d;
                return (int)((sw[word_of(d)] >> (((D >> 2) & 1) * 16)) & 0xffffu);
            };
            short res;
            // first minimum in internal order = minimum of the keys sad * 64 + d
            // (each word holds D = kk + 8 jj in its low half, D + 4 in its high half)
            unsigned pk[4];                                      // one chain per kk: short dependency chains
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                pk[kk] = 0xffffffffu;
#pragma unroll
                for (int jj = 0; jj < DN_NJ / 2; ++jj) {
                    const unsigned w = sw[(kk * (DN_NJ / 2) + jj) * DN_TW];
                    const unsigned d_lo = DN_ND - 1 - (kk + 8 * jj), d_hi = d_lo - 4;
                    pk[kk] = min(pk[kk], min(((w << 6) & 0x3fffc0u) | d_lo, ((w >> 10) & 0x3fffc0u) | d_hi));
                }
            }
            const unsigned ukey = min(min(pk[0], pk[1]), min(pk[2], pk[3]));
            const int key = (int)ukey;
            const int mind = key & 63, minsad = key >> 6;
            bool ok = s_tex[k * DN_TW + lane] >= texture;
            const int pp = sad_at(mind + 1 < DN_ND ? mind + 1 : DN_ND - 2);
            const int nn = sad_at(mind > 0 ? mind - 1 : 1);
            if (uniq > 0) {
                // rejected iff some d outside mind-1 .. mind+1 has sad <= thresh: knock those
                // three out (0xffff) and take the packed 16-bit minimum of the 24 words
                const int thresh = minsad + (minsad * uniq / 100);
#pragma unroll
                for (int t = -1; t <= 1; ++t) {
                    const int d = mind + t;
                    if (d >= 0 && d < DN_ND)
                        sw[word_of(d)] |= 0xffffu << ((((DN_ND - 1 - d) >> 2) & 1) * 16);
                }
                unsigned pm[4];
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    pm[kk] = 0xffffffffu;
#pragma unroll
                    for (int jj = 0; jj < DN_NJ / 2; ++jj)
                        pm[kk] = __vminu2(pm[kk], sw[(kk * (DN_NJ / 2) + jj) * DN_TW]);
                }
                const unsigned m = __vminu2(__vminu2(pm[0], pm[1]), __vminu2(pm[2], pm[3]));
                if ((int)min(m & 0xffffu, m >> 16) <= thresh) ok = false;
            }
            if (ok) {
                const int c0 = minsad;
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
        // tall bands amortise the 20 start-up rows of a band (KITTI: 2 bands of 180 rows); short
        // ones give a lone image enough blocks for 148 SMs x 2
        const int tiles = (vw + DN_TW - 1) / DN_TW;
        const int nb_tall = (vh + DN_BAND_TALL - 1) / DN_BAND_TALL;
        int band = DN_BAND;
        if ((long long)tiles * nb_tall * n_call >= 2 * 296)       // two waves of 2 blocks x 148 SMs
            band = ((vh + nb_tall - 1) / nb_tall + 3) & ~3;       // equal bands, <= DN_BAND_TALL
        const size_t smem = sizeof(unsigned) * (DN_RING_WORDS + DN_SAD_WORDS + DN_TEXR_WORDS + DN_TEX_WORDS) +
                            (size_t)(band + DN_WIN - 1) * (DN_PLW + DN_PRW);
        // per device, so not cached: costs a microsecond
        SVO_CUDA_OK(cudaFuncSetAttribute(stereo_bm_dense_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dim3 grid(tiles, (vh + band - 1) / band, n_call);
        stereo_bm_dense_kernel<<<grid, 128, smem, st>>>(pyr, lay, seq_stride, right, right_stride,
                                                        disp, dpitch, disp_stride, band, sp.cap,
                                                        sp.texture, sp.uniq, d_call);
    }
    return cudaGetLastError();
}
