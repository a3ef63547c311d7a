// lk_track: cv::calcOpticalFlowPyrLK(prev, next, pts, Size(21,21), maxLevel 3,
// {COUNT+EPS, 30, 0.01}, flags 0, minEigThreshold 1e-2) - reference call sites
// src/feature_tracker.cpp:23-26 (forward) and :32-35 (backward); arithmetic pinned in
// oracle/restate.py:lk_track.
//
// One block of 64 threads per key-point, walking pyramid levels top -> 0 like LKTrackerInvoker.
// Thread (wy, part) owns pixels [0,8), [8,16) or [16,21) of window row wy (21 rows x 3 parts =
// 63 threads) - the two 8-pixel SIMD steps and the 5-pixel scalar tail of OpenCV's row loop -
// so neighbouring taps are shared in registers and the pixel pairs (x, x+4) OpenCV multiplies
// together in one v_dotprod lane sit in one thread.
//   * the 24x24 raw patch of I is staged in shared memory: with aligned 32-bit loads when it
//     lies inside the image (the common case), byte-wise with reflect-101 at the border;
//   * interior windows: B = w00*I + w01*I(x+1) + w10*I(y+1) + w11*I(x+1,y+1) is built once on a
//     23x23 grid (exact integers). The interpolated patch is (B + 2^8) >> 9 and, because the
//     unnormalised Scharr operator is linear, the interpolated derivatives OpenCV takes from
//     its int16 derivative image are exactly (Scharr(B) + 2^13) >> 14 - no derivative image is
//     ever materialised;
//   * windows touching the image border keep OpenCV's semantics literally: Scharr of the
//     reflect-101 image on the 22x22 integer grid, ZERO outside the image, then interpolation;
//   * A11/A12/A22 and, per iteration, b1/b2 are accumulated IN OPENCV'S ORDER, in f32 with a
//     rounding after every addition: four SIMD lanes fed by pixels (L, L+4, L+8, L+12) of every
//     row, a scalar tail fed by pixels 16..20, lanes folded as (l0+l2)+(l1+l3), tail added last
//     (b: int32 sums of the pixel pairs (x, x+4), converted, two accumulators of four lanes).
//     The per-pixel terms go through shared memory in chain order and ten (b) / fifteen (A)
//     lanes of warp 0 each run one sequential chain. With exact integer sums instead the
//     result differed from cv2 in the last bits of ~20 % of the points (<= 2e-4 px) - enough to
//     flip the 0.01-px termination test once in ~1e6 iterations, after which a track drifts
//     from the reference's by up to 0.05 px for the rest of its life. Now flow is bit-exact
//     (oracle/restate.py:lk_window_sums pins the order against cv2);
//   * a 28x28 region of J is staged per level (aligned words when interior) and re-staged only
//     when the window leaves it, so an iteration is shared-memory + integer work only; all
//     threads carry the scalar Gauss-Newton state redundantly: uniform control flow.
//   * shared memory is the busiest unit (ncu: 70 % of its wavefront rate, more than half of
//     them bank-conflict replays with byte loads and 32-byte rows), so staged rows are 36 bytes
//     apart, a thread fetches its 8 J bytes of two rows as 6 aligned words (instead of 16 byte
//     loads) and blends two taps per dp2a: -14 % kernel time, same integers.
// Issue/shared-memory bound (<= 400 points x 4 levels x <= 30 iterations), not HBM-bound.
#include "svo_common.cuh"

namespace {

constexpr int WIN = SVO_LK_WIN;            // 21
constexpr int NT = 64;                     // threads per key-point
constexpr int RUN = 8;                     // window pixels per thread (the tail part uses 5)
constexpr int IP = WIN + 3;                // raw I patch edge (24)
constexpr int SP = 36;                     // byte pitch of the staged I / J rows (9 words:
                                           // rows 4 apart no longer share banks)
constexpr int BG = WIN + 2;                // B grid edge (23)
constexpr int BP = 26;                     // int pitch of the B grid
constexpr int DG = WIN + 1;                // derivative grid edge, border path (22)
constexpr int JM = 3;                      // margin of the staged J region
constexpr int JR = WIN + 1 + 2 * JM;       // staged J region edge (28)
constexpr unsigned FULL = 0xffffffffu;
constexpr int LK_LEVELS = 4;               // levels pre-staged (reference: maxLevel 3)

// ---- window sums in OpenCV's accumulation order (lkpyramid.cpp, CV_SIMD128 path) -----------
// Terms are laid out in shared memory by CHAIN: a chain is one f32 accumulator of the OpenCV
// loop, its terms in the order that loop adds them.
//   A (per level): chain q*4+L, q in {A11, A12, A22}, SIMD lane L: 84 terms, term (y, m) = pixel
//   (y, L + 4m); then the three scalar tails of 105 terms, term (y, t) = pixel (y, 16 + t).
//   b (per iteration): chain p*2+xy, p = pixel pair (p, p+4) of a SIMD step, xy in {b1, b2}: 42
//   terms, term (y, blk) = f32(int32 pair sum) of step blk; then the two tails of 105 terms.
constexpr int A_LANE_TERMS = 4 * WIN;                      // 84
constexpr int TAIL_TERMS = 5 * WIN;                        // 105
constexpr int TAIL_PITCH = 108;                            // float4-aligned
constexpr int A_TAIL_OFF = 12 * A_LANE_TERMS;              // 1008
constexpr int A_FLOATS = A_TAIL_OFF + 3 * TAIL_PITCH;      // 1332
constexpr int B_CHAIN_TERMS = 2 * WIN;                     // 42
constexpr int B_CHAIN_PITCH = 44;
// b layout: the four b1 chains, then the four b2 chains B_YOFF floats further on; the two tails
// behind them, again B_YOFF apart - so a thread's b2 term always sits B_YOFF behind its b1 term
// and one piece of code (base + p * stride) serves the SIMD parts and the tail part
constexpr int B_YOFF = 4 * B_CHAIN_PITCH;                  // 176
constexpr int B_TAIL_OFF = 2 * B_YOFF;                     // 352
constexpr int B_DUMMY_OFF = B_TAIL_OFF + 2 * B_YOFF;       // 704: where the idle thread 63 writes
static_assert(TAIL_PITCH <= B_YOFF, "tail chains do not overlap");
static_assert(B_DUMMY_OFF + B_YOFF + 3 * B_CHAIN_PITCH + 5 <= A_FLOATS, "b terms fit in the A term buffer");
// control word of the next iteration (s_ctl.z): >= 0 byte offset of the window origin in the
// staged J region, or one of these
constexpr int CTL_STOP = -1;               // the iteration loop of this level is over
constexpr int CTL_RESTAGE = -2;            // the window left the staged region: re-stage at .w

// acc = fl(acc + term) over the terms at src (16-byte aligned), in order: N_SHORT terms for a
// SIMD-lane chain, TAIL_TERMS for a scalar-tail chain (is_long); inactive lanes idle. The
// dependent additions are the critical path (TAIL_TERMS x the FADD latency per window sum).
// Loops of five / seven 4-term groups instead of straight-line code: the kernel's hot code has
// to stay inside the 32 KB instruction cache of the SM (ncu: a fifth of the warp stalls were
// instruction fetches when both chains were fully unrolled, 2.4 KB each).
__device__ __forceinline__ float add4(float acc, const float* p) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    return __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, v.x), v.y), v.z), v.w);
}
template <int N_SHORT>
__device__ __forceinline__ float chain_sum(const float* __restrict__ src, bool active, bool is_long) {
    constexpr int G_ALL = N_SHORT / 4, REM = N_SHORT % 4;     // 10, 2 (b) / 21, 0 (A)
    constexpr int G_TAIL = TAIL_TERMS / 4;                     // 26 full groups, then one term
    constexpr int G_MID = G_ALL + (REM ? 1 : 0);
    static_assert(TAIL_TERMS % 4 == 1 && (G_ALL % 5 == 0 || G_ALL % 7 == 0), "group loops");
    float acc = 0.f;
    if (active) {
#pragma unroll(G_ALL % 5 == 0 ? 5 : 7)
        for (int g = 0; g < G_ALL; ++g) acc = add4(acc, src + 4 * g);
        if (REM) {
            const float4 v = *reinterpret_cast<const float4*>(src + 4 * G_ALL);
            acc = __fadd_rn(acc, v.x);
            if (REM > 1 || is_long) acc = __fadd_rn(acc, v.y);
            if (REM > 2 || is_long) acc = __fadd_rn(acc, v.z);
            if (is_long) acc = __fadd_rn(acc, v.w);
        }
    }
    if (is_long) {
#pragma unroll 5
        for (int g = G_MID; g < G_TAIL; ++g) acc = add4(acc, src + 4 * g);
        acc = __fadd_rn(acc, src[4 * G_TAIL]);
    }
    return acc;
}

// c + w.lo16 * p.byte0 + w.hi16 * p.byte1 (lo) / ... byte2, byte3 (hi): signed 16-bit weights
// (w11 can be -1 after rounding), unsigned pixels
__device__ __forceinline__ int dp2a_lo(int w, unsigned p, int c) {
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(p), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi(int w, unsigned p, int c) {
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(p), "r"(c));
    return d;
}
// Which of the two warps runs the window-sum chains and the scalar step of iteration j (j = -1:
// plans the first iteration) of level l for key-point pi.
//   0: always warp 0   1: fixed per block   2: changes with the level   3: with every iteration
#ifndef LK_CW_MODE
#define LK_CW_MODE 3
#endif
__device__ __forceinline__ int chain_warp(int j, int l, int pi) {
    if (LK_CW_MODE == 0) return 0;
    if (LK_CW_MODE == 1) return (pi >> 1) & 1;
    if (LK_CW_MODE == 2) return ((pi >> 1) + l) & 1;
    return ((pi >> 1) + l + j + 1) & 1;
}

__device__ __forceinline__ int pack16(int lo, int hi) { return (lo & 0xffff) | (hi << 16); }

// bytes [k, k+1] of the 8-byte run (a0 | a1 << 32) as the low two bytes of a word
template <int K>
__device__ __forceinline__ int tap2(int w, unsigned a0, unsigned a1, int c) {
    if (K % 4 == 0) return dp2a_lo(w, K < 4 ? a0 : a1, c);
    if (K % 4 == 2) return dp2a_hi(w, K < 4 ? a0 : a1, c);
    if (K == 3) return dp2a_lo(w, __funnelshift_r(a0, a1, 24), c);
    return dp2a_lo(w, (K < 4 ? a0 : a1) >> 8, c);          // K % 4 == 1 (K == 7: see tap2_7)
}
// bytes [7, 8] of the run: the top byte of a1 and the low byte of the next word
__device__ __forceinline__ int tap2_7(int w, unsigned a1, unsigned a2, int c) {
    return dp2a_lo(w, __funnelshift_r(a1, a2, 24), c);
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

// Staging of an N_ROWS x N_COLS byte region whose first wanted byte is (x0, y0) into shared
// rows of SP bytes. Column range inside the image (the common case, also for windows that
// stick out above/below: rows are fetched through reflect-101 indices): aligned words copied
// global -> shared by cp.async (no registers in between, so the regions of all levels are in
// flight together and the code stays small: thread = word column tid & 7 of rows tid >> 3,
// + 8, + 16, ...). Returns x0 & 3 (byte offset of x0 in a staged row). Column range crossing the
// image border: byte-wise reflect-101, done at once; returns 0. The caller waits
// (stage_wait) and puts a barrier before anybody reads.
__device__ __forceinline__ void cp_async4(unsigned dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void stage_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int N_ROWS, int N_COLS>
__device__ __forceinline__ int stage_region(const uint8_t* __restrict__ img, int w, int h,
                                            int pitch, int x0, int y0, unsigned* s_words,
                                            int tid) {
    constexpr int NW = (N_COLS + 3 + 3) / 4;          // words covering any alignment
    static_assert(NW <= 8 && NW * 4 <= SP, "staged row too wide");
    if (x0 >= 0 && x0 + N_COLS <= w) {
        const int off = x0 & 3;
        const int col = tid & 7, row0 = tid >> 3;
        if (col < NW) {
            const uint8_t* base = img + (x0 - off) + col * 4;
            const unsigned dst = (unsigned)__cvta_generic_to_shared(s_words) + row0 * SP + col * 4;
#pragma unroll
            for (int p = 0; p < (N_ROWS + 7) / 8; ++p) {
                const int r = row0 + 8 * p;
                if (r < N_ROWS)
                    cp_async4(dst + p * 8 * SP, base + (size_t)clamp_reflect(y0 + r, h) * pitch);
            }
        }
        return off;
    }
    unsigned char* s_bytes = reinterpret_cast<unsigned char*>(s_words);
    for (int i = tid; i < N_ROWS * N_COLS; i += NT) {
        int r = i / N_COLS, c = i - r * N_COLS;
        int gy = clamp_reflect(y0 + r, h), gx = clamp_reflect(x0 + c, w);
        s_bytes[r * SP + c] = img[(size_t)gy * pitch + gx];
    }
    return 0;
}

__global__ void __launch_bounds__(NT, 16)
lk_track_kernel(const uint8_t* __restrict__ pyr, const __grid_constant__ SvoPyrLayout lay,
                size_t seq_stride,
                const float2* __restrict__ in_pts, size_t in_stride, int in_buffered,
                float2* __restrict__ out_pts, unsigned char* __restrict__ status, int out_stride,
                int direction, int max_iters, double eps2, double min_eig_thr, int max_pts,
                const int* __restrict__ gate, int gate_min,
                const unsigned char* __restrict__ skip_status,
                const SvoSeqCall* __restrict__ call, const SvoLkCache cache) {
    // I patch and J region of EVERY level are fetched up front (their positions are known
    // before any iteration: pt / 2^l, zero-flow guess for J), so a point pays the global-memory
    // latency once instead of once or twice per level; J is re-staged only when the window
    // leaves the +-JM margin of its level's region
    __shared__ __align__(16) unsigned s_Iall[LK_LEVELS][IP * (SP / 4)];
    __shared__ __align__(16) unsigned s_Jall[LK_LEVELS][JR * (SP / 4)];
    // B grid of the interior path (border path: s_dx / s_dy) and, once every thread has taken
    // its window values from it, the chain-ordered terms of the window sums
    __shared__ __align__(16) float s_T[A_FLOATS];
    __shared__ float s_sum[3];
    __shared__ int4 s_ctl;
    __shared__ float4 s_next;              // nextPt and ok when a level's iterations end
    __shared__ float4 s_st;                // np, prevDelta between iterations (LK_CW_MODE 3)
    int* s_B = reinterpret_cast<int*>(s_T);
    static_assert(sizeof(int) * BG * BP <= sizeof(float) * A_FLOATS, "B grid fits in the term buffer");

    const SvoSeqCall* cp = call + blockIdx.y;
    const int seq = cp->seq;
    const int pi = blockIdx.x;
    if (pi >= cp->n_feat || (cp->flags & 1) == 0) return;
    if (gate && gate[seq] < gate_min) return;      // < 4 corners: the reference returns first
    const int tid = threadIdx.x;
    const int wrp = tid >> 5;

    const float2 pt = in_pts[(size_t)seq * in_stride +
                             (in_buffered ? (size_t)cp->feat_buf * max_pts : 0) + pi];
    if (skip_status && skip_status[(size_t)seq * out_stride + pi] == 0) {
        // backward pass of a point the forward pass lost: the reference tracks it anyway
        // (feature_tracker.cpp:32-35) but never looks at the result (:45-47)
        if (tid == 0) {
            out_pts[(size_t)seq * out_stride + pi] = pt;
            status[(size_t)seq * out_stride + pi] = 0;
        }
        return;
    }

    const int cur_slot = cp->cur_slot;
    const uint8_t* seq_base = pyr + (size_t)seq * seq_stride;
    const int slot_i = direction == 0 ? (1 - cur_slot) : cur_slot;
    const uint8_t* __restrict__ pyrI = seq_base + (size_t)slot_i * lay.slot_bytes;
    const uint8_t* __restrict__ pyrJ = seq_base + (size_t)(1 - slot_i) * lay.slot_bytes;

    // Hand-over between passes. The forward pass of frame t+1 looks at the same image (frame t)
    // around the same points (the tracked positions) as the backward pass of frame t did: all
    // that depends on I alone - the interpolated window I, Ix, Iy of every level, A11/A12/A22,
    // 1/D, the min-eigenvalue verdict - is computed once, by the backward pass, and parked in
    // global memory (12 KB per point as int16, OpenCV's own IWinBuf type). The forward pass
    // takes it if the entry its feature came from carries the serial of this pyramid build and
    // this very point, bit for bit; otherwise (new features, feature sets installed by the host,
    // first frames) it computes everything as before. Either way the numbers are the same.
    const bool c_store = direction == 1 && cache.data != nullptr;
    bool cached = false;
    unsigned c_status = 0;
    size_t c_entry = (size_t)seq * max_pts + pi;
    unsigned c_serial = 0;
    if (cache.data != nullptr) {
        c_serial = cache.img_serial[seq * 2 + slot_i];
        if (direction == 0) {
            const int src = cache.kept_idx[(size_t)seq * max_pts + pi];
            c_entry = (size_t)seq * max_pts + min(max(src, 0), max_pts - 1);
            const uint4 hd = __ldcg(reinterpret_cast<const uint4*>(cache.hdr + c_entry));
            cached = c_serial != 0u && hd.x == c_serial && hd.y == __float_as_uint(pt.x) &&
                     hd.z == __float_as_uint(pt.y);
            c_status = hd.w;
            // every level whose window origin lies inside the bordered image must be there
            for (int l = 0; cached && l < lay.n_levels; ++l) {
                const SvoLevel lv = lay.lvl[l];
                const float inv = 1.f / (float)(1 << l);
                const int ipx = (int)floorf(__fsub_rn(__fmul_rn(pt.x, inv), 10.f));
                const int ipy = (int)floorf(__fsub_rn(__fmul_rn(pt.y, inv), 10.f));
                const unsigned st = (c_status >> (8 * l)) & 0xffu;
                if (!(ipx < -WIN || ipx >= lv.w || ipy < -WIN || ipy >= lv.h) && st != 1u && st != 3u)
                    cached = false;
            }
        }
    }

    // window pixels owned by this thread: row wy, columns wx0 .. wx0 + 7 (part 2: wx0 + 4;
    // thread 63 idles)
    const bool has_px = tid < WIN * 3;
    const int wy = min(tid / 3, WIN - 1);
    const int part = tid - (tid / 3) * 3;
    const int wx0 = part * RUN;
    const int lane = tid & 31;
    // where this thread's b terms go: term p at s_T[t_off + p * t_str] (b1) and B_YOFF behind it
    // (b2); the idle thread 63 writes to a spare area
    const int t_off = tid >= WIN * 3 ? B_DUMMY_OFF : part < 2 ? wy * 2 + part : B_TAIL_OFF + wy * 5;
    const int t_str = part < 2 ? B_CHAIN_PITCH : 1;
    const int tb = wy * SP + wx0;          // byte offset of its first window pixel in a staged region
    short* s_dx = reinterpret_cast<short*>(s_B);
    short* s_dy = s_dx + DG * DG;
    static_assert(2 * DG * DG * sizeof(short) <= sizeof(int) * BG * BP, "alias too small");

    const float half = 10.f;                 // (winSize - 1) * 0.5
    const float FLT_SCALE = 1.f / (float)(1 << 20);

    // ---- fetch the I patch and the zero-flow J region of every level (loads of all levels in
    // flight together), then one barrier
    unsigned offs = 0;                       // 2 bits per (level, I/J): byte offset in a staged row
#pragma unroll 1
    for (int l = 0; l < lay.n_levels; ++l) {
        const SvoLevel lv = lay.lvl[l];
        const float inv = 1.f / (float)(1 << l);
        const int ipx = (int)floorf(__fsub_rn(__fmul_rn(pt.x, inv), half));
        const int ipy = (int)floorf(__fsub_rn(__fmul_rn(pt.y, inv), half));
        if (!(ipx < -WIN || ipx >= lv.w || ipy < -WIN || ipy >= lv.h)) {
            int oi = 0;
            if (!cached)
                oi = stage_region<IP, IP>(pyrI + lv.off, lv.w, lv.h, lv.pitch, ipx - 1, ipy - 1,
                                          s_Iall[l], tid);
            const int oj = stage_region<JR, JR>(pyrJ + lv.off, lv.w, lv.h, lv.pitch, ipx - JM,
                                                ipy - JM, s_Jall[l], tid);
            offs |= (unsigned)(oi | (oj << 2)) << (4 * l);
        }
    }
    stage_wait();
    __syncthreads();
    unsigned st_word = 0;                    // c_store: one status byte per level
    float2 nextPt = make_float2(0.f, 0.f);   // the value OpenCV keeps in nextPts[ptidx]
    bool ok = true;

    for (int l = lay.n_levels - 1; l >= 0; --l) {
        const SvoLevel lv = lay.lvl[l];
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
        const int offI = (offs >> (4 * l)) & 3;      // staged before the loop
        const unsigned char* s_Ib = reinterpret_cast<const unsigned char*>(s_Iall[l]);
        unsigned* s_Jw = s_Jall[l];
        __syncthreads();                     // previous level's readers are done with s_B

        int ival[RUN], ixv[RUN], iyv[RUN];
        float A11, A12, A22, D;
        if (cached) {
            // the backward pass of the previous frame left this level's window state
            if (((c_status >> (8 * l)) & 0xffu) != 1u) {      // min-eigenvalue reject
                if (l == 0) ok = false;
                continue;
            }
            const uint4* cd = cache.data + ((c_entry * LK_LEVELS + l) * 3) * NT + tid;
            const uint4 q0 = __ldcs(cd), q1 = __ldcs(cd + NT), q2 = __ldcs(cd + 2 * NT);
            const float4 ca = __ldcg(&cache.hdr[c_entry].a[l]);
            const unsigned wi[4] = {q0.x, q0.y, q0.z, q0.w}, wx[4] = {q1.x, q1.y, q1.z, q1.w},
                           wv[4] = {q2.x, q2.y, q2.z, q2.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                ival[2 * k] = 256 - (int)((wi[k] & 0xffffu) << 9);
                ival[2 * k + 1] = 256 - (int)((wi[k] >> 16) << 9);
                ixv[2 * k] = (int)(short)(wx[k] & 0xffffu);
                ixv[2 * k + 1] = (int)wx[k] >> 16;
                iyv[2 * k] = (int)(short)(wv[k] & 0xffffu);
                iyv[2 * k + 1] = (int)wv[k] >> 16;
            }
            A11 = ca.x; A12 = ca.y; A22 = ca.z; D = ca.w;
        } else {
        int w00, w01, w10, w11;
        lk_weights(__fsub_rn(prevPt.x, (float)ipx), __fsub_rn(prevPt.y, (float)ipy), w00, w01,
                   w10, w11);
        // every corner of every window pixel inside the image <=> no zero-border derivative
        const bool grid_inside = ipx >= 0 && ipy >= 0 && ipx + WIN < w && ipy + WIN < h;
        if (grid_inside) {
            if (tid < 2 * BG) {              // B grid: thread = (row, half row)
                // 13 bytes of two staged rows as aligned words (4 per row), blended two taps
                // at a time by dp2a: exact integers, a quarter of the shared-memory requests
                const int r = tid >> 1, c0 = (tid & 1) * 12, n = (tid & 1) ? 11 : 12;
                const int bo = r * SP + offI + c0;
                const unsigned* q0 = reinterpret_cast<const unsigned*>(s_Ib) + (bo >> 2);
                const unsigned* q1 = q0 + SP / 4;
                const unsigned sh = (bo & 3) * 8;
                // rows are SP = 36 bytes apart and 24 + 3 bytes are staged: words 0..3 of the
                // second half stay inside the row's 9 words (c0 = 12: word index <= 3 + 4)
                const unsigned u0 = q0[0], u1 = q0[1], u2 = q0[2], u3 = q0[3];
                const unsigned v0 = q1[0], v1 = q1[1], v2 = q1[2], v3 = q1[3];
                const unsigned a0 = __funnelshift_r(u0, u1, sh), a1 = __funnelshift_r(u1, u2, sh),
                               a2 = __funnelshift_r(u2, u3, sh), a3 = u3 >> sh;
                const unsigned b0 = __funnelshift_r(v0, v1, sh), b1 = __funnelshift_r(v1, v2, sh),
                               b2 = __funnelshift_r(v2, v3, sh), b3 = v3 >> sh;
                const int wt = pack16(w00, w01), wb = pack16(w10, w11);
                int* o = s_B + r * BP + c0;
                int ov[12];
                ov[0] = tap2<0>(wb, b0, b1, tap2<0>(wt, a0, a1, 0));
                ov[1] = tap2<1>(wb, b0, b1, tap2<1>(wt, a0, a1, 0));
                ov[2] = tap2<2>(wb, b0, b1, tap2<2>(wt, a0, a1, 0));
                ov[3] = tap2<3>(wb, b0, b1, tap2<3>(wt, a0, a1, 0));
                ov[4] = tap2<0>(wb, b1, b2, tap2<0>(wt, a1, a2, 0));
                ov[5] = tap2<1>(wb, b1, b2, tap2<1>(wt, a1, a2, 0));
                ov[6] = tap2<2>(wb, b1, b2, tap2<2>(wt, a1, a2, 0));
                ov[7] = tap2<3>(wb, b1, b2, tap2<3>(wt, a1, a2, 0));
                ov[8] = tap2<0>(wb, b2, b3, tap2<0>(wt, a2, a3, 0));
                ov[9] = tap2<1>(wb, b2, b3, tap2<1>(wt, a2, a3, 0));
                ov[10] = tap2<2>(wb, b2, b3, tap2<2>(wt, a2, a3, 0));
                ov[11] = tap2<3>(wb, b2, b3, tap2<3>(wt, a2, a3, 0));
#pragma unroll
                for (int k = 0; k < 12; ++k)
                    if (k < n) o[k] = ov[k];
            }
            __syncthreads();
            const int* bp = s_B + wy * BP + wx0;
            int t0[RUN + 2], t1[RUN + 2];
#pragma unroll
            for (int c = 0; c < RUN + 2; ++c) {
                int b0 = bp[c], b1 = bp[BP + c], b2 = bp[2 * BP + c];
                t0[c] = 3 * (b0 + b2) + 10 * b1;
                t1[c] = b2 - b0;
                if (c >= 1 && c <= RUN) ival[c - 1] = (b1 + 256) >> 9;
            }
#pragma unroll
            for (int t = 0; t < RUN; ++t) {
                ixv[t] = (t0[t + 2] - t0[t] + 8192) >> 14;
                iyv[t] = (3 * (t1[t] + t1[t + 2]) + 10 * t1[t + 1] + 8192) >> 14;
            }
        } else {
            // border path: derivative image semantics of OpenCV (zero outside the image)
            for (int i = tid; i < DG * DG; i += NT) {
                int r = i / DG, c = i - r * DG;
                int gy = ipy + r, gx = ipx + c;
                int vx = 0, vy = 0;
                if (gx >= 0 && gx < w && gy >= 0 && gy < h) {
                    const unsigned char* q = s_Ib + r * SP + offI + c;   // pixel (gy-1, gx-1)
                    int a0 = q[0], a1 = q[1], a2 = q[2];
                    int b0 = q[SP], b2 = q[SP + 2];
                    int c0 = q[2 * SP], c1 = q[2 * SP + 1], c2 = q[2 * SP + 2];
                    int t0l = 3 * (a0 + c0) + 10 * b0, t0r = 3 * (a2 + c2) + 10 * b2;
                    int t1l = c0 - a0, t1c = c1 - a1, t1r = c2 - a2;
                    vx = t0r - t0l;
                    vy = 3 * (t1l + t1r) + 10 * t1c;
                }
                s_dx[i] = (short)vx;
                s_dy[i] = (short)vy;
            }
            __syncthreads();
#pragma unroll
            for (int t = 0; t < RUN; ++t) {
                const int wx = min(wx0 + t, WIN - 1);      // columns >= WIN: computed, never used
                const unsigned char* q = s_Ib + (wy + 1) * SP + offI + wx + 1;
                ival[t] = (q[0] * w00 + q[1] * w01 + q[SP] * w10 + q[SP + 1] * w11 + 256) >> 9;
                const short* dx = s_dx + wy * DG + wx;
                const short* dy = s_dy + wy * DG + wx;
                ixv[t] = (dx[0] * w00 + dx[1] * w01 + dx[DG] * w10 + dx[DG + 1] * w11 + 8192) >> 14;
                iyv[t] = (dy[0] * w00 + dy[1] * w01 + dy[DG] * w10 + dy[DG + 1] * w11 + 8192) >> 14;
            }
        }
        if (c_store) {                       // hand the window to the next frame's forward pass
            unsigned wi[4], wx[4], wv[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                wi[k] = __byte_perm((unsigned)ival[2 * k], (unsigned)ival[2 * k + 1], 0x5410);
                wx[k] = __byte_perm((unsigned)ixv[2 * k], (unsigned)ixv[2 * k + 1], 0x5410);
                wv[k] = __byte_perm((unsigned)iyv[2 * k], (unsigned)iyv[2 * k + 1], 0x5410);
            }
            uint4* cd = cache.data + ((c_entry * LK_LEVELS + l) * 3) * NT + tid;
            __stcs(cd, make_uint4(wi[0], wi[1], wi[2], wi[3]));
            __stcs(cd + NT, make_uint4(wx[0], wx[1], wx[2], wx[3]));
            __stcs(cd + 2 * NT, make_uint4(wv[0], wv[1], wv[2], wv[3]));
        }
#pragma unroll
        for (int t = 0; t < RUN; ++t) ival[t] = 256 - (ival[t] << 9);   // see the iteration
        // ---- A11, A12, A22: per-pixel f32 products in chain order, then sequential chains.
        // Every thread forms the products of its 8 pixels (one piece of code for the SIMD parts
        // and the tail part, whose last three are never stored); chain (q, L) takes pixels L and
        // L + 4 of a SIMD part as two ADJACENT terms: one 64-bit store
        __syncthreads();                     // every thread has its values out of s_B / s_dx
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            float pr[RUN];
#pragma unroll
            for (int t = 0; t < RUN; ++t)
                pr[t] = __int2float_rn((q == 2 ? iyv[t] : ixv[t]) * (q == 0 ? ixv[t] : iyv[t]));
            if (part < 2) {
                if (has_px) {
                    float2* o = reinterpret_cast<float2*>(s_T + q * 4 * A_LANE_TERMS + wy * 4 + 2 * part);
#pragma unroll
                    for (int L = 0; L < 4; ++L) o[L * (A_LANE_TERMS / 2)] = make_float2(pr[L], pr[L + 4]);
                }
            } else {
                float* o = s_T + A_TAIL_OFF + q * TAIL_PITCH + wy * 5;
#pragma unroll
                for (int t = 0; t < WIN - 2 * RUN; ++t) o[t] = pr[t];
            }
        }
        __syncthreads();
        if (wrp == chain_warp(-2, l, pi)) {
            const float* src = lane < 12 ? s_T + lane * A_LANE_TERMS
                                         : s_T + A_TAIL_OFF + (min(lane, 14) - 12) * TAIL_PITCH;
            const float acc = chain_sum<A_LANE_TERMS>(src, lane < 15, lane >= 12 && lane < 15);
            // v_reduce_sum: (l0 + l2) + (l1 + l3), then the scalar tail takes it in
            const float u = __fadd_rn(acc, __shfl_down_sync(FULL, acc, 2));
            const float red = __fadd_rn(u, __shfl_down_sync(FULL, u, 1));
            const float tail = __shfl_sync(FULL, acc, 12 + (lane >> 2));
            if (lane < 12 && (lane & 3) == 0) s_sum[lane >> 2] = __fadd_rn(tail, red);
        }
        __syncthreads();
        const float iA11 = s_sum[0], iA12 = s_sum[1], iA22 = s_sum[2];
        A11 = __fmul_rn(iA11, FLT_SCALE);
        A12 = __fmul_rn(iA12, FLT_SCALE);
        A22 = __fmul_rn(iA22, FLT_SCALE);
        D = __fsub_rn(__fmul_rn(A11, A22), __fmul_rn(A12, A12));
        const float dif = __fsub_rn(A11, A22);
        const float rad = __fadd_rn(__fmul_rn(dif, dif), __fmul_rn(__fmul_rn(4.f, A12), A12));
        const float minEig =
            __fdiv_rn(__fsub_rn(__fadd_rn(A22, A11), __fsqrt_rn(rad)), (float)(2 * WIN * WIN));
        if ((double)minEig < min_eig_thr || D < 1.1920928955078125e-7f) {
            st_word |= 3u << (8 * l);
            if (l == 0) ok = false;
            continue;
        }
        D = __fdiv_rn(1.f, D);
        if (c_store) {
            st_word |= 1u << (8 * l);
            if (tid == 0) cache.hdr[c_entry].a[l] = make_float4(A11, A12, A22, D);
        }
        }

        // b terms: a SIMD part sums the pixel pairs (p, p + 4) in int32 (OpenCV's v_dotprod
        // lanes), the tail part takes pixels 0..4 one by one. With the upper derivatives of the
        // tail part zeroed, "pair p" is the same expression for both: term p of this thread
        int ixh[4], iyh[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            ixh[p] = part < 2 ? ixv[p + 4] : 0;
            iyh[p] = part < 2 ? iyv[p + 4] : 0;
        }
        const int ix4 = ixv[4], iy4 = iyv[4];        // the tail part's fifth pixel

        // Scalar Gauss-Newton state lives in ONE warp at a time (the warp that also runs the
        // chains): it plans every iteration - window origin, bounds test, bilinear weights,
        // position in the staged J region - and hands the plan to the block in one 16-byte
        // control word. Which warp that is changes (chain_warp): the chain warp issues three
        // times the instructions of the other one, and the two warps of a block sit on
        // different schedulers of the SM.
        float2 np = make_float2(__fsub_rn(nextPt.x, half), __fsub_rn(nextPt.y, half));
        float2 prevDelta = make_float2(0.f, 0.f);
        int jx0 = ipx - JM, jy0 = ipy - JM;          // origin of the pre-staged J region
        int offJ = (offs >> (4 * l + 2)) & 3;
        auto plan = [&]() {                          // one warp, every lane the same values
            int4 c = make_int4(0, 0, CTL_STOP, 0);
            const int inx = (int)floorf(np.x), iny = (int)floorf(np.y);
            if (inx < -WIN || inx >= w || iny < -WIN || iny >= h) {
                if (l == 0) ok = false;
                if (lane == 0) s_next = make_float4(nextPt.x, nextPt.y, ok ? 1.f : 0.f, 0.f);
            } else {
                int v00, v01, v10, v11;
                lk_weights(__fsub_rn(np.x, (float)inx), __fsub_rn(np.y, (float)iny), v00, v01,
                           v10, v11);
                c.x = pack16(v00, v01);
                c.y = pack16(v10, v11);
                const int ox = inx - jx0, oy = iny - jy0;
                if (ox < 0 || ox > 2 * JM || oy < 0 || oy > 2 * JM) {
                    c.z = CTL_RESTAGE;
                    c.w = (inx & 0xffff) | (iny << 16);
                } else {
                    c.z = oy * SP + ox;
                }
                if (LK_CW_MODE == 3 && lane == 0)
                    s_st = make_float4(np.x, np.y, prevDelta.x, prevDelta.y);
            }
            if (lane == 0) s_ctl = c;
        };
        if (wrp == chain_warp(-1, l, pi)) {
            if (max_iters > 0) plan();
            else if (lane == 0) {
                s_next = make_float4(nextPt.x, nextPt.y, ok ? 1.f : 0.f, 0.f);
                s_ctl = make_int4(0, 0, CTL_STOP, 0);
            }
        }
        __syncthreads();
        for (int j = 0;; ++j) {                      // every thread counts the iterations
            const int4 ctl = s_ctl;
            int z = ctl.z;
            if (z == CTL_STOP) break;                // uniform: every thread reads the same word
            if (z == CTL_RESTAGE) {
                // (the last readers of the region passed two barriers ago)
                const int sx = (int)(short)(ctl.w & 0xffff), sy = ctl.w >> 16;
                jx0 = sx - JM;
                jy0 = sy - JM;
                offJ = stage_region<JR, JR>(J, w, h, pitch, jx0, jy0, s_Jw, tid);
                stage_wait();
                __syncthreads();
                z = JM * SP + JM;
            }
            {
                // the thread's 9 bytes of two J rows as aligned words + funnel shifts; a pixel's
                // four taps are two dp2a (exact integers): 6 word loads instead of 18 byte loads
                const int bo = tb + offJ + z;
                const unsigned* q0 = s_Jw + (bo >> 2);
                const unsigned* q1 = q0 + SP / 4;
                const unsigned sh = (bo & 3) * 8;
                const unsigned u0 = q0[0], u1 = q0[1], u2 = q0[2];
                const unsigned v0 = q1[0], v1 = q1[1], v2 = q1[2];
                const unsigned a0 = __funnelshift_r(u0, u1, sh), a1 = __funnelshift_r(u1, u2, sh),
                               a2 = u2 >> sh;
                const unsigned b0 = __funnelshift_r(v0, v1, sh), b1 = __funnelshift_r(v1, v2, sh),
                               b2 = v2 >> sh;
                const int wt = ctl.x, wb = ctl.y;
                // ((J + 256) >> 9) - I = (J + 256 - (I << 9)) >> 9: the I term rides in the dp2a
                // accumulator (ival holds 256 - (I << 9))
                int dI[RUN];
                dI[0] = tap2<0>(wb, b0, b1, tap2<0>(wt, a0, a1, ival[0])) >> 9;
                dI[1] = tap2<1>(wb, b0, b1, tap2<1>(wt, a0, a1, ival[1])) >> 9;
                dI[2] = tap2<2>(wb, b0, b1, tap2<2>(wt, a0, a1, ival[2])) >> 9;
                dI[3] = tap2<3>(wb, b0, b1, tap2<3>(wt, a0, a1, ival[3])) >> 9;
                dI[4] = tap2<4>(wb, b0, b1, tap2<4>(wt, a0, a1, ival[4])) >> 9;
                dI[5] = tap2<5>(wb, b0, b1, tap2<5>(wt, a0, a1, ival[5])) >> 9;
                dI[6] = tap2<6>(wb, b0, b1, tap2<6>(wt, a0, a1, ival[6])) >> 9;
                dI[7] = tap2_7(wb, b1, b2, tap2_7(wt, a1, a2, ival[7])) >> 9;
                float* o = s_T + t_off;
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    o[p * t_str] = __int2float_rn(dI[p] * ixv[p] + dI[p + 4] * ixh[p]);
                    o[p * t_str + B_YOFF] = __int2float_rn(dI[p] * iyv[p] + dI[p + 4] * iyh[p]);
                }
                if (part == 2) {
                    o[4] = __int2float_rn(dI[4] * ix4);
                    o[4 + B_YOFF] = __int2float_rn(dI[4] * iy4);
                }
            }
            __syncthreads();
            if (wrp == chain_warp(j, l, pi)) {
                if (LK_CW_MODE == 3) {               // the state the planner left
                    const float4 st = s_st;
                    np = make_float2(st.x, st.y);
                    prevDelta = make_float2(st.z, st.w);
                }
                // chain `lane` = pair p = lane >> 1 of b1 (even lanes) / b2 (odd lanes); lanes
                // 8, 9: the scalar tails
                const float* src = lane < 8 ? s_T + (lane & 1) * B_YOFF + (lane >> 1) * B_CHAIN_PITCH
                                            : s_T + B_TAIL_OFF + (min(lane, 9) - 8) * B_YOFF;
                const float acc = chain_sum<B_CHAIN_TERMS>(src, lane < 10, lane >= 8 && lane < 10);
                // qb0 + qb1 (chains p and p + 2), then lanes (0, 2) -> b1, (1, 3) -> b2, then the
                // scalar tails take the SIMD sums in
                const float q = __fadd_rn(acc, __shfl_down_sync(FULL, acc, 4));
                const float r = __fadd_rn(q, __shfl_down_sync(FULL, q, 2));
                const float tail = __shfl_sync(FULL, acc, 8 + (lane & 1));
                const float tot = __fadd_rn(tail, r);
                const float ib1 = __shfl_sync(FULL, tot, 0), ib2 = __shfl_sync(FULL, tot, 1);
                const float b1 = __fmul_rn(ib1, FLT_SCALE);
                const float b2 = __fmul_rn(ib2, FLT_SCALE);
                float2 delta;
                delta.x = __fmul_rn(__fsub_rn(__fmul_rn(A12, b2), __fmul_rn(A22, b1)), D);
                delta.y = __fmul_rn(__fsub_rn(__fmul_rn(A12, b1), __fmul_rn(A11, b2)), D);
                np.x = __fadd_rn(np.x, delta.x);
                np.y = __fadd_rn(np.y, delta.y);
                nextPt = make_float2(__fadd_rn(np.x, half), __fadd_rn(np.y, half));
                const double dd = (double)delta.x * (double)delta.x + (double)delta.y * (double)delta.y;
                bool stop = dd <= eps2;
                if (!stop && j > 0 && fabs((double)__fadd_rn(delta.x, prevDelta.x)) < 0.01 &&
                    fabs((double)__fadd_rn(delta.y, prevDelta.y)) < 0.01) {
                    nextPt.x = __fsub_rn(nextPt.x, __fmul_rn(delta.x, 0.5f));
                    nextPt.y = __fsub_rn(nextPt.y, __fmul_rn(delta.y, 0.5f));
                    stop = true;
                }
                prevDelta = delta;
                if (j + 1 >= max_iters) stop = true;
                if (!stop) plan();
                else if (lane == 0) {
                    s_next = make_float4(nextPt.x, nextPt.y, ok ? 1.f : 0.f, 0.f);
                    s_ctl = make_int4(0, 0, CTL_STOP, 0);
                }
            }
            __syncthreads();
        }
        {   // the chain warp's result: where the next level starts, whether the point is alive
            const float4 nx = s_next;
            nextPt = make_float2(nx.x, nx.y);
            ok = nx.z != 0.f;
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
        out_pts[(size_t)seq * out_stride + pi] = nextPt;
        status[(size_t)seq * out_stride + pi] = ok ? 1 : 0;
        // (counted here, not at the look-up: two more live values up there made ptxas spill in
        // the level prologue and cost the kernel 17 %)
        if (direction == 0 && cache.stats) {
            atomicAdd(cache.stats, 1ull);
            if (cached) atomicAdd(cache.stats + 1, 1ull);
        }
        if (c_store)                         // the entry is complete: key it
            *reinterpret_cast<uint4*>(cache.hdr + c_entry) =
                make_uint4(c_serial, __float_as_uint(pt.x), __float_as_uint(pt.y), st_word);
    }
}

// ---- hand-over after the host installed a feature set ---------------------------------------
// svo_tracker_set replaces the tracker's feature list (inliers picked by the host's pose step +
// new features, feature_tracker.cpp:3-16), so the filter's index map no longer leads from
// feature k to the entry the backward pass wrote for it. This kernel rebuilds the map: feature
// k points at the entry that carries the current image's serial and the feature's very bits
// (the tracked positions travel to the host and back unchanged), or at entry 0 when there is
// none (new features) - the look-up in lk_track checks the key either way, so the map can only
// turn recomputation into a hit, never change a number. It runs beside the copy of the list,
// not inside lk_track: a scan there cost registers lk_track does not have.
constexpr int RK_T = 256;
__global__ void __launch_bounds__(RK_T)
lk_rekey_kernel(const float2* __restrict__ feat, int max_pts,
                const __grid_constant__ SvoLkRekey rk, SvoLkCacheHdr* hdr,
                const unsigned int* __restrict__ img_serial, int* __restrict__ kept_idx) {
    __shared__ unsigned int s_x[RK_T], s_y[RK_T], s_ok[RK_T];
    const int ci = blockIdx.y;
    const int seq = rk.seq[ci];
    const int count = rk.count[ci];
    const int k = blockIdx.x * RK_T + threadIdx.x;
    if (blockIdx.x * RK_T >= count) return;              // whole block idle (uniform)
    const unsigned int serial = img_serial[seq * 2 + rk.slot[ci]];
    unsigned int px = 0u, py = 0u;
    if (k < count) {
        const float2 p = feat[((size_t)seq * 2 + rk.feat_buf[ci]) * max_pts + k];
        px = __float_as_uint(p.x);
        py = __float_as_uint(p.y);
    }
    int found = 0;
    for (int j0 = 0; j0 < max_pts; j0 += RK_T) {
        const int j = j0 + threadIdx.x;
        unsigned int ok = 0u, hx = 0u, hy = 0u;
        if (j < max_pts) {
            const uint4 hd = __ldcg(reinterpret_cast<const uint4*>(hdr + (size_t)seq * max_pts + j));
            ok = (serial != 0u && hd.x == serial) ? 1u : 0u;
            hx = hd.y;
            hy = hd.z;
        }
        __syncthreads();
        s_x[threadIdx.x] = hx;
        s_y[threadIdx.x] = hy;
        s_ok[threadIdx.x] = ok;
        __syncthreads();
        const int m = min(RK_T, max_pts - j0);
        for (int t = 0; t < m; ++t)
            if (s_ok[t] && s_x[t] == px && s_y[t] == py) found = j0 + t;
    }
    if (k < count) kept_idx[(size_t)seq * max_pts + k] = found;
}

}  // namespace

cudaError_t svo_launch_lk_rekey(const float2* feat, int max_pts,
                                const SvoLkRekey& rk, const SvoLkCache& cache, int* kept_idx,
                                cudaStream_t st) {
    int most = 0;
    for (int i = 0; i < rk.n; ++i) most = rk.count[i] > most ? rk.count[i] : most;
    if (rk.n <= 0 || most <= 0 || cache.hdr == nullptr || max_pts <= 0) return cudaSuccess;
    dim3 grid((most + RK_T - 1) / RK_T, rk.n);
    lk_rekey_kernel<<<grid, RK_T, 0, st>>>(feat, max_pts, rk, cache.hdr,
                                           cache.img_serial, kept_idx);
    return cudaGetLastError();
}

cudaError_t svo_launch_lk_track(const uint8_t* pyr, const SvoPyrLayout& lay, size_t seq_stride,
                                const float2* in_pts, size_t in_stride, int in_buffered,
                                float2* out_pts, unsigned char* status, int out_stride,
                                int direction, SvoLkParams lp, int max_pts, const int* gate,
                                int gate_min, const unsigned char* skip_status,
                                const SvoSeqCall* d_call, int n_call, cudaStream_t st,
                                SvoLkCache cache) {
    if (max_pts <= 0 || n_call <= 0) return cudaSuccess;
    // 16 blocks of 64 threads and 13.9 KB per SM: the window-sum chains are latency-bound
    // (one dependent FADD every 4 cycles), so resident blocks are what fills the issue slots
    static const cudaError_t carve = cudaFuncSetAttribute(
        lk_track_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (carve != cudaSuccess) return carve;
    dim3 grid(lp.grid_pts > 0 ? lp.grid_pts : max_pts, n_call);
    lk_track_kernel<<<grid, NT, 0, st>>>(pyr, lay, seq_stride, in_pts, in_stride, in_buffered,
                                         out_pts, status, out_stride, direction, lp.max_iters,
                                         lp.eps2, lp.min_eig, max_pts, gate, gate_min,
                                         skip_status, d_call, cache);
    return cudaGetLastError();
}
