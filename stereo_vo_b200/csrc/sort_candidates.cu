// sort_candidates: step 8 of cv::goodFeaturesToTrack (reference call site
// src/image_processor.cpp:22): order the NMS survivors by score descending, ties by larger
// linear address first. Keys are (monotone score key << 32 | address), so one descending
// 64-bit sort gives exactly that order; keys are unique, hence the result is deterministic.
//
// Bitonic network. Up to 16384 candidates per sequence (KITTI: 2k-11k) are sorted by one block
// entirely in shared memory (128 KB); larger capacities (4K stress config) use 8192-key tiles
// in shared memory plus global compare-exchange passes for the strides that span tiles.
// Latency-bound, not HBM-bound: n*log2(n)^2/2 compare-exchanges on on-chip data.
#include "svo_common.cuh"

namespace {

constexpr int SORT_THREADS = 1024;
constexpr int SORT_SMEM_MAX = 16384;       // keys sorted by a single block
constexpr int SORT_TILE = 8192;            // tile of the multi-block path

__device__ __forceinline__ int next_pow2(int n) {
    int p = 2;
    while (p < n) p <<= 1;
    return p;
}

// one bitonic compare-exchange step on shared memory; `base` is the global index of s[0]
__device__ __forceinline__ void bitonic_step_smem(unsigned long long* s, int count, int base,
                                                  int k, int j) {
    for (int t = threadIdx.x; t < (count >> 1); t += blockDim.x) {
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int l = i + j;
        unsigned long long a = s[i], b = s[l];
        bool desc = (((base + i) & k) == 0);
        if ((a < b) == desc) { s[i] = b; s[l] = a; }
    }
}

// All remaining steps of a merge phase once the stride is <= 32: the 32 pairs of consecutive
// pair indices touch one block of 64 consecutive keys for every such stride, so a warp runs
// them back to back on its own blocks with warp barriers only (a block barrier per step is what
// dominates the sort of a few thousand keys).
__device__ __forceinline__ void bitonic_tail_smem(unsigned long long* s, int count, int base,
                                                  int k, int j_first) {
    const int lane = threadIdx.x & 31;
    for (int t0 = (threadIdx.x >> 5) << 5; t0 < (count >> 1); t0 += blockDim.x) {
        const int t = t0 + lane;
        for (int j = j_first; j > 0; j >>= 1) {
            if (t < (count >> 1)) {
                int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                int l = i + j;
                unsigned long long a = s[i], b = s[l];
                bool desc = (((base + i) & k) == 0);
                if ((a < b) == desc) { s[i] = b; s[l] = a; }
            }
            __syncwarp();
        }
    }
}

// Single-block path: gather the per-strip candidate lists of corner_score_nms, apply the exact
// quality threshold (step 6: thr = (float)((double)max * quality), keep eig > thr), sort.
__global__ void __launch_bounds__(SORT_THREADS)
sort_small_kernel(unsigned long long* __restrict__ cand, int max_cand,
                  int* __restrict__ cand_count, const unsigned long long* __restrict__ strips,
                  int n_strips, int strip_cap, const int* __restrict__ strip_count,
                  const unsigned int* __restrict__ eig_max, double quality,
                  const SvoSeqCall* __restrict__ call) {
    extern __shared__ unsigned long long s_keys[];
    __shared__ int s_n, s_over;
    const int seq = call[blockIdx.x].seq;
    unsigned long long* keys = cand + (size_t)seq * max_cand;
    if (threadIdx.x == 0) { s_n = 0; s_over = 0; }
    __syncthreads();
    const unsigned mk = eig_max[seq];
    const float maxv = mk ? __uint_as_float(svo_key_float_bits(mk)) : 0.f;
    const float thr = (float)((double)maxv * quality);
    const int lane = threadIdx.x & 31;
    // one warp per candidate list (there are n_strips x n_seg short ones): the count -> list
    // load chains of different lists overlap, and a warp keeps 8 loads per lane in flight
    constexpr int GB = 8;
    for (int st = threadIdx.x >> 5; st < n_strips; st += SORT_THREADS / 32) {
        const int raw = strip_count[seq * n_strips + st];
        const int cnt = min(raw, strip_cap);
        if (raw > strip_cap && lane == 0) s_over = 1;
        const unsigned long long* src = strips + ((size_t)seq * n_strips + st) * (size_t)strip_cap;
        for (int i0 = 0; i0 < cnt; i0 += 32 * GB) {
            unsigned long long key[GB];
#pragma unroll
            for (int c = 0; c < GB; ++c) {
                const int i = i0 + c * 32 + lane;
                key[c] = (i < cnt) ? src[i] : 0ull;
            }
            unsigned keep_mask = 0;
#pragma unroll
            for (int c = 0; c < GB; ++c) {
                const float e = __uint_as_float(svo_key_float_bits((unsigned)(key[c] >> 32)));
                // padding keys are 0: score key 0 decodes to a NaN pattern, never > thr
                if ((i0 + c * 32 + lane < cnt) && (e > thr) && (e != 0.f)) keep_mask |= 1u << c;
            }
            int mine = __popc(keep_mask);
            // exclusive prefix of `mine` over the warp, total in every lane
            int incl = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += v;
            }
            const int total_w = __shfl_sync(0xffffffffu, incl, 31);
            if (total_w == 0) continue;
            int base = 0;
            if (lane == 0) base = atomicAdd(&s_n, total_w);
            base = __shfl_sync(0xffffffffu, base, 0);
            int idx = base + incl - mine;
#pragma unroll
            for (int c = 0; c < GB; ++c) {
                if (keep_mask & (1u << c)) {
                    if (idx < max_cand) s_keys[idx] = key[c];
                    ++idx;
                }
            }
        }
    }
    __syncthreads();
    const int total = s_n;
    if (threadIdx.x == 0) cand_count[seq] = s_over ? (1 << 30) : total;   // > max_cand: overflow
    const int n = min(total, max_cand);
    if (n == 0) return;
    const int npad = next_pow2(n);
    for (int i = n + threadIdx.x; i < npad; i += blockDim.x) s_keys[i] = 0ull;
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1) {
        int j = k >> 1;
        for (; j >= 64; j >>= 1) {
            bitonic_step_smem(s_keys, npad, 0, k, j);
            __syncthreads();
        }
        bitonic_tail_smem(s_keys, npad, 0, k, j);
        __syncthreads();
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) keys[i] = s_keys[i];
}

// ---- multi-block path -------------------------------------------------------------------
// phase A: every tile fully sorted (k = 2 .. SORT_TILE), zero padding written back
__global__ void __launch_bounds__(SORT_THREADS)
sort_tile_kernel(unsigned long long* __restrict__ cand, int cap_pad,
                 const int* __restrict__ cand_count, int max_cand,
                 const SvoSeqCall* __restrict__ call) {
    extern __shared__ unsigned long long s_keys[];
    const int seq = call[blockIdx.y].seq;
    unsigned long long* keys = cand + (size_t)seq * cap_pad;
    const int n = min(cand_count[seq], max_cand);
    const int npad = max(next_pow2(n), SORT_TILE);
    const int base = blockIdx.x * SORT_TILE;
    if (base >= npad) return;
    for (int i = threadIdx.x; i < SORT_TILE; i += blockDim.x)
        s_keys[i] = (base + i < n) ? keys[base + i] : 0ull;
    __syncthreads();
    for (int k = 2; k <= SORT_TILE; k <<= 1) {
        int j = k >> 1;
        for (; j >= 64; j >>= 1) {
            bitonic_step_smem(s_keys, SORT_TILE, base, k, j);
            __syncthreads();
        }
        bitonic_tail_smem(s_keys, SORT_TILE, base, k, j);
        __syncthreads();
    }
    for (int i = threadIdx.x; i < SORT_TILE; i += blockDim.x) keys[base + i] = s_keys[i];
}

// phase B: one compare-exchange pass with stride j >= SORT_TILE
__global__ void __launch_bounds__(256)
sort_global_step_kernel(unsigned long long* __restrict__ cand, int cap_pad,
                        const int* __restrict__ cand_count, int max_cand, int k, int j,
                        const SvoSeqCall* __restrict__ call) {
    const int seq = call[blockIdx.y].seq;
    unsigned long long* keys = cand + (size_t)seq * cap_pad;
    const int n = min(cand_count[seq], max_cand);
    const int npad = max(next_pow2(n), SORT_TILE);
    if (k > npad) return;
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (npad >> 1)) return;
    int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
    int l = i + j;
    unsigned long long a = keys[i], b = keys[l];
    bool desc = ((i & k) == 0);
    if ((a < b) == desc) { keys[i] = b; keys[l] = a; }
}

// phase C: finish merge level k inside each tile (j = SORT_TILE/2 .. 1)
__global__ void __launch_bounds__(SORT_THREADS)
sort_tile_merge_kernel(unsigned long long* __restrict__ cand, int cap_pad,
                       const int* __restrict__ cand_count, int max_cand, int k,
                       const SvoSeqCall* __restrict__ call) {
    extern __shared__ unsigned long long s_keys[];
    const int seq = call[blockIdx.y].seq;
    unsigned long long* keys = cand + (size_t)seq * cap_pad;
    const int n = min(cand_count[seq], max_cand);
    const int npad = max(next_pow2(n), SORT_TILE);
    const int base = blockIdx.x * SORT_TILE;
    if (k > npad || base >= npad) return;
    for (int i = threadIdx.x; i < SORT_TILE; i += blockDim.x) s_keys[i] = keys[base + i];
    __syncthreads();
    for (int j = SORT_TILE >> 1; j >= 64; j >>= 1) {
        bitonic_step_smem(s_keys, SORT_TILE, base, k, j);
        __syncthreads();
    }
    bitonic_tail_smem(s_keys, SORT_TILE, base, k, 32);
    __syncthreads();
    for (int i = threadIdx.x; i < SORT_TILE; i += blockDim.x) keys[base + i] = s_keys[i];
}

}  // namespace

// `max_cand` is the logical capacity; the buffer stride is the capacity padded to a power of
// two (svo_sort_capacity_pad) when the multi-block path is used, and max_cand otherwise.
int svo_sort_capacity_pad(int max_cand) {
    if (max_cand <= SORT_SMEM_MAX) return max_cand;
    int p = SORT_TILE;
    while (p < max_cand) p <<= 1;
    return p;
}

int svo_sort_launch_count(int max_cand) {
    if (max_cand <= SORT_SMEM_MAX) return 1;
    const int cap_pad = svo_sort_capacity_pad(max_cand);
    int c = 2;                                  // filter + tile sort
    for (int k = SORT_TILE * 2; k <= cap_pad; k <<= 1) {
        for (int j = k >> 1; j >= SORT_TILE; j >>= 1) ++c;
        ++c;
    }
    return c;
}

// opt in to > 48 KB dynamic shared memory on the current device (call once per device)
cudaError_t svo_sort_init() {
    SVO_CUDA_OK(cudaFuncSetAttribute(sort_small_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     SORT_SMEM_MAX * 8));
    SVO_CUDA_OK(cudaFuncSetAttribute(sort_tile_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, SORT_TILE * 8));
    SVO_CUDA_OK(cudaFuncSetAttribute(sort_tile_merge_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, SORT_TILE * 8));
    return cudaSuccess;
}

cudaError_t svo_launch_sort_candidates(unsigned long long* cand, int max_cand, int* cand_count,
                                       const unsigned long long* strips, int n_strips,
                                       int strip_cap, const int* strip_count,
                                       const unsigned int* eig_max, double quality,
                                       const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    if (max_cand <= SORT_SMEM_MAX) {
        int p = 2;
        while (p < max_cand) p <<= 1;
        sort_small_kernel<<<n_call, SORT_THREADS, (size_t)p * 8, st>>>(
            cand, max_cand, cand_count, strips, n_strips, strip_cap, strip_count, eig_max,
            quality, d_call);
        return cudaGetLastError();
    }
    const int cap_pad = svo_sort_capacity_pad(max_cand);
    const int tiles = cap_pad / SORT_TILE;
    SVO_CUDA_OK(svo_launch_filter_candidates(strips, n_strips, strip_cap, strip_count, eig_max,
                                             quality, cand, cap_pad, max_cand, cand_count, d_call,
                                             n_call, st));
    sort_tile_kernel<<<dim3(tiles, n_call), SORT_THREADS, SORT_TILE * 8, st>>>(
        cand, cap_pad, cand_count, max_cand, d_call);
    for (int k = SORT_TILE * 2; k <= cap_pad; k <<= 1) {
        for (int j = k >> 1; j >= SORT_TILE; j >>= 1) {
            sort_global_step_kernel<<<dim3((cap_pad / 2 + 255) / 256, n_call), 256, 0, st>>>(
                cand, cap_pad, cand_count, max_cand, k, j, d_call);
        }
        sort_tile_merge_kernel<<<dim3(tiles, n_call), SORT_THREADS, SORT_TILE * 8, st>>>(
            cand, cap_pad, cand_count, max_cand, k, d_call);
    }
    return cudaGetLastError();
}
