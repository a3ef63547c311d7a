// greedy_min_distance: steps 9-10 of cv::goodFeaturesToTrack (reference call site
// src/image_processor.cpp:22; pinned in oracle/restate.py:greedy_min_distance).
// Sequential semantics: walk candidates in sorted order, accept one iff no already accepted
// corner lies closer than minDistance, stop after maxCorners.
//
// Exact parallel form, one block per sequence: candidates are taken in chunks in rank order (the
// first 256, then 1024 at a time: the head of the list is where most corners are accepted, the
// long tail is almost entirely rejected by step 1 and only costs barriers). (1) every thread
// tests its candidate against the corners accepted by earlier chunks (cell grid, 3x3 cells,
// <= 4 corners per cell); (2) survivors are compacted in rank order - at most 256 per chunk: a
// chunk with more ends at its 257th survivor, which opens the next chunk - and
// their pairwise conflicts go into 256-bit masks; (3) one warp resolves the chunk sequentially
// (survivor b is accepted iff none of its conflicting predecessors was accepted); (4) accepted
// corners are appended in rank order and inserted into the grid. Output is identical to the
// sequential algorithm because acceptance of a candidate depends only on higher-ranked ones.
// Latency-bound (on-chip), not HBM-bound.
#include "svo_common.cuh"

namespace {

constexpr int GS_THREADS = 1024;           // candidates examined per chunk (after the first)
constexpr int GS_HEAD = 256;               // size of the first chunk
constexpr int GS_SURV = 256;               // survivors resolved per chunk
constexpr int GS_WORDS = GS_SURV / 32;
constexpr int GS_WARPS = GS_THREADS / 32;
constexpr int CELL_CAP = 4;
constexpr int CELL_INTS = CELL_CAP + 1;    // count + packed (x | y << 16)
constexpr unsigned FULL = 0xffffffffu;

__global__ void __launch_bounds__(GS_THREADS)
greedy_select_kernel(const unsigned long long* __restrict__ cand, int max_cand,
                     const int* __restrict__ cand_count, int W, int H, double min_distance,
                     int max_corners, float2* __restrict__ corners, int* __restrict__ n_corners,
                     int* grid_scratch, size_t grid_stride, int use_smem_grid,
                     const SvoSeqCall* __restrict__ call) {
    extern __shared__ int s_grid_dyn[];
    __shared__ int s_sx[GS_SURV], s_sy[GS_SURV];
    __shared__ unsigned s_conf[GS_SURV][GS_WORDS];
    __shared__ unsigned s_acc[GS_WORDS];
    __shared__ int s_wcnt[GS_WARPS];
    __shared__ int s_next;                     // candidate index of the first survivor left out

    const int seq = call[blockIdx.x].seq;
    const unsigned long long* keys = cand + (size_t)seq * max_cand;
    float2* out = corners + (size_t)seq * max_corners;
    const int n = min(cand_count[seq], max_cand);
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;

    if (min_distance < 1.0) {                  // OpenCV skips the spacing rule entirely
        int m = min(n, max_corners);
        for (int i = tid; i < m; i += GS_THREADS) {
            unsigned addr = (unsigned)(keys[i] & 0xffffffffull);
            int y = addr / W, x = addr - y * W;
            out[i] = make_float2((float)x, (float)y);
        }
        if (tid == 0) n_corners[seq] = m;
        return;
    }

    const int cell = (int)rint(min_distance);  // cvRound
    const int gw = (W + cell - 1) / cell, gh = (H + cell - 1) / cell;
    // dx*dx + dy*dy < minDistance^2 on integers: for an integer d2, d2 < t <=> d2 < ceil(t)
    const double md2 = min_distance * min_distance;
    const int imd2 = (int)fmin(ceil(md2), 2147483647.0);
    int* grid = use_smem_grid ? s_grid_dyn : (grid_scratch + (size_t)seq * grid_stride);
    for (int i = tid; i < gw * gh; i += GS_THREADS) grid[i * CELL_INTS] = 0;
    __syncthreads();

    int accepted_total = 0;
    for (int c0 = 0; c0 < n && accepted_total < max_corners;) {
        const int chunk = (c0 == 0) ? GS_HEAD : GS_THREADS;
        const int i = c0 + tid;
        int x = 0, y = 0;
        bool alive = false;
        if (tid == 0) s_next = 0x7fffffff;
        if (tid < chunk && i < n) {
            unsigned addr = (unsigned)(keys[i] & 0xffffffffull);
            y = addr / W;
            x = addr - y * W;
            alive = true;
            const int xc = x / cell, yc = y / cell;
            const int x1 = max(0, xc - 1), x2 = min(gw - 1, xc + 1);
            const int y1 = max(0, yc - 1), y2 = min(gh - 1, yc + 1);
            for (int yy = y1; yy <= y2 && alive; ++yy)
                for (int xx = x1; xx <= x2 && alive; ++xx) {
                    const int* cp = grid + (yy * gw + xx) * CELL_INTS;
                    int cnt = min(cp[0], CELL_CAP);
                    for (int k = 0; k < cnt; ++k) {
                        int p = cp[1 + k];
                        int dx = x - (p & 0xffff), dy = y - (p >> 16);
                        if (dx * dx + dy * dy < imd2) { alive = false; break; }
                    }
                }
        }
        // ---- rank-ordered compaction of the survivors (the first GS_SURV of them)
        unsigned bal = __ballot_sync(FULL, alive);
        if (lane == 0) s_wcnt[wrp] = __popc(bal);
        __syncthreads();
        int off = 0, m_all = 0;
#pragma unroll
        for (int w = 0; w < GS_WARPS; ++w) {
            int c = s_wcnt[w];
            if (w < wrp) off += c;
            m_all += c;
        }
        if (alive) {
            int pos = off + __popc(bal & ((1u << lane) - 1u));
            if (pos < GS_SURV) {
                s_sx[pos] = x;
                s_sy[pos] = y;
            } else if (pos == GS_SURV) {
                s_next = i;                    // exactly one thread: the 257th survivor
            }
        }
        __syncthreads();
        const int m = min(m_all, GS_SURV);
        // candidates behind the last resolved survivor wait for the next chunk (those among
        // them that step 1 rejected stay rejected: the accepted set only grows)
        const int c_next = (m_all > GS_SURV) ? s_next : c0 + chunk;
        if (m == 0) { c0 = c_next; __syncthreads(); continue; }     // uniform
        // ---- pairwise conflicts with earlier survivors of this chunk
        if (tid < m) {
            unsigned msk[GS_WORDS];
#pragma unroll
            for (int w = 0; w < GS_WORDS; ++w) msk[w] = 0;
            const int bx = s_sx[tid], by = s_sy[tid];
            for (int a = 0; a < tid; ++a) {
                int dx = bx - s_sx[a], dy = by - s_sy[a];
                if (dx * dx + dy * dy < imd2) msk[a >> 5] |= 1u << (a & 31);
            }
#pragma unroll
            for (int w = 0; w < GS_WORDS; ++w) s_conf[tid][w] = msk[w];
        }
        __syncthreads();
        // ---- sequential resolution by warp 0 (lane w < 8 owns accepted-mask word w)
        if (wrp == 0) {
            unsigned acc = 0;
            for (int b = 0; b < m; ++b) {
                unsigned c = (lane < GS_WORDS) ? (s_conf[b][lane] & acc) : 0u;
                unsigned any = __ballot_sync(FULL, c != 0);
                if (!any && lane == (b >> 5)) acc |= 1u << (b & 31);
            }
            if (lane < GS_WORDS) s_acc[lane] = acc;
        }
        __syncthreads();
        // ---- append accepted corners in rank order, insert into the grid
        int chunk_acc = 0, rank = 0;
#pragma unroll
        for (int w = 0; w < GS_WORDS; ++w) {
            int c = __popc(s_acc[w]);
            if (w < (tid >> 5)) rank += c;
            chunk_acc += c;
        }
        if (tid < m) {
            unsigned word = s_acc[tid >> 5];
            if ((word >> (tid & 31)) & 1u) {
                rank += __popc(word & ((1u << (tid & 31)) - 1u));
                int o = accepted_total + rank;
                if (o < max_corners) {
                    const int px = s_sx[tid], py = s_sy[tid];
                    out[o] = make_float2((float)px, (float)py);
                    int* cp = grid + ((py / cell) * gw + (px / cell)) * CELL_INTS;
                    int slot = atomicAdd(cp, 1);
                    if (slot < CELL_CAP) cp[1 + slot] = px | (py << 16);
                }
            }
        }
        accepted_total += chunk_acc;
        c0 = c_next;
        __syncthreads();
    }
    if (tid == 0) n_corners[seq] = min(accepted_total, max_corners);
}

}  // namespace

int svo_greedy_grid_ints(int w, int h, float min_distance) {
    if (min_distance < 1.f) return 0;
    int cell = (int)rint((double)min_distance);
    int gw = (w + cell - 1) / cell, gh = (h + cell - 1) / cell;
    return gw * gh * CELL_INTS;
}

cudaError_t svo_greedy_init() {
    return cudaFuncSetAttribute(greedy_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                96 * 1024);
}

cudaError_t svo_launch_greedy_select(const unsigned long long* cand, int max_cand,
                                     const int* cand_count, int w, int h, float min_distance,
                                     int max_corners, float2* corners, int* n_corners,
                                     int* grid_scratch, size_t grid_stride,
                                     const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    size_t grid_bytes = (size_t)svo_greedy_grid_ints(w, h, min_distance) * sizeof(int);
    int use_smem = grid_bytes <= 96 * 1024;
    greedy_select_kernel<<<n_call, GS_THREADS, use_smem ? grid_bytes : 0, st>>>(
        cand, max_cand, cand_count, w, h, (double)min_distance, max_corners, corners, n_corners,
        grid_scratch, grid_stride, use_smem, d_call);
    return cudaGetLastError();
}
