// svo_dedupe_static: the new-vs-tracked dedupe of svo_dedupe (stereo.cu: dedupe_kernel; reference
// rule src/image_processor.cpp:113-128) for callers that must not allocate: no engine, no
// cudaMalloc / cudaFree / cudaHostAlloc, nothing but a memset node and one kernel on the
// caller's stream, so it can sit inside a CUDA graph capture or a serving loop. NO COUNTERPART IN
// THE REFERENCE (include/svo_c.h says the same).
//
// The kernel is dedupe_kernel's arithmetic for ONE point set: blocks test their share of the
// detections against the tracked points (staged in shared memory, same fp32 operation order,
// same threshold on the squared distance) and leave a keep flag per detection in scratch; the
// block that takes the last ticket compacts in input order. Scratch = the ticket word + one
// flag per detection. In svo_dedupe that scratch belongs to an engine (d_ticket, d_keep_idx);
// here it is the file-scope array below: (64 + 16384) ints = 65 792 bytes of device memory on
// every GPU that loads libsvo.so, held until the process exits.
//
// Every call on a device shares that one array. The call clears the ticket itself (the flags
// are written before they are read), so a call does not depend on what the previous one left;
// but two calls that OVERLAP on the GPU (different streams, or graph branches without an edge
// between them) race on it. Calls on one stream, or ordered by events, are safe.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/svo_c.h"

namespace {

constexpr int DS_THREADS = 256;
constexpr int DS_TILE = 1024;              // tracked points staged per pass
constexpr int DS_MAX_BLOCKS = 32;
constexpr int DS_HEADER = 64;              // ints in front of the flags: [0] = ticket
constexpr int DS_MAX_DET = SVO_DEDUPE_STATIC_MAX_DET;
constexpr int DS_MAX_DEVICES = 64;
constexpr unsigned FULL = 0xffffffffu;

// Device memory. In HOST code this name is only the symbol's host-side handle, never an
// address a kernel may be given: ds_scratch_of() asks the runtime for the device address.
static __device__ __align__(256) int g_dedupe_scratch[DS_HEADER + DS_MAX_DET];

// smallest float t with sqrt_rn(t) >= md: "sqrt_rn(d2) < md" as a test on d2 (stereo.cu has
// the same function for dedupe_kernel)
__device__ __forceinline__ float ds_dist2_threshold(float md) {
    if (!(md > 0.f)) return 0.f;
    float t = __fmul_rn(md, md);
    for (int k = 0; k < 8 && t > 0.f; ++k) {
        const float lo = __int_as_float(__float_as_int(t) - 1);
        if (__fsqrt_rn(lo) >= md) t = lo; else break;
    }
    for (int k = 0; k < 8 && __fsqrt_rn(t) < md; ++k) t = __int_as_float(__float_as_int(t) + 1);
    return t;
}

__global__ void __launch_bounds__(DS_THREADS)
dedupe_static_kernel(const float2* __restrict__ det, int nd, const float2* __restrict__ trk,
                     int ni, float min_distance, float2* __restrict__ out,
                     int* __restrict__ keep_idx, int* __restrict__ n_out, int* scratch) {
    __shared__ float2 s_q[DS_TILE];
    __shared__ int s_w[DS_THREADS / 32];
    __shared__ bool s_last;
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    int* flag = scratch + DS_HEADER;
    const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
    const float thr = ds_dist2_threshold(min_distance);
    const int per = (nd + (int)gridDim.x - 1) / (int)gridDim.x;
    const int i_beg = blockIdx.x * per, i_end = min(nd, i_beg + per);
    for (int i0 = i_beg; i0 < i_end; i0 += DS_THREADS) {
        const int i = i0 + tid;
        const float2 p = (i < i_end) ? det[i] : make_float2(0.f, 0.f);
        bool near = false;
        for (int j0 = 0; j0 < ni; j0 += DS_TILE) {
            const int nj = min(DS_TILE, ni - j0);
            __syncthreads();
            for (int j = tid; j < nj; j += DS_THREADS) s_q[j] = trk[j0 + j];
            __syncthreads();
#pragma unroll 8
            for (int j = 0; j < nj; ++j) {
                const float2 t = s_q[j];
                const float dx = __fsub_rn(p.x, t.x), dy = __fsub_rn(p.y, t.y);
                near |= __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < thr;
            }
        }
        if (i < i_end) flag[i] = near ? 0 : 1;
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    int written = 0;
    for (int i0 = 0; i0 < nd; i0 += DS_THREADS) {
        const int i = i0 + tid;
        const bool keep = i < nd && __ldcg(flag + i) != 0;
        const unsigned bal = __ballot_sync(FULL, keep);
        if (lane == 0) s_w[wrp] = __popc(bal);
        __syncthreads();
        int off = 0, total = 0;
        for (int w = 0; w < DS_THREADS / 32; ++w) {
            const int c = s_w[w];
            if (w < wrp) off += c;
            total += c;
        }
        __syncthreads();
        if (keep) {
            const int slot = written + off + __popc(bal & ((1u << lane) - 1u));
            if (out) out[slot] = det[i];
            if (keep_idx) keep_idx[slot] = i;
        }
        written += total;
    }
    if (tid == 0) {
        *n_out = written;
        *ticket = 0;
    }
}

// device address of the scratch array, looked up once per device with that device current
std::mutex g_ds_mu;
int* g_ds_ptr[DS_MAX_DEVICES];

cudaError_t ds_scratch_of(int device, int** out) {
    std::lock_guard<std::mutex> lk(g_ds_mu);
    if (!g_ds_ptr[device]) {
        void* p = nullptr;
        cudaError_t ce = cudaGetSymbolAddress(&p, g_dedupe_scratch);
        if (ce != cudaSuccess) return ce;
        g_ds_ptr[device] = static_cast<int*>(p);
    }
    *out = g_ds_ptr[device];
    return cudaSuccess;
}

int ds_fail(char* err, size_t cap, int code, const char* what, cudaError_t ce = cudaSuccess) {
    if (err && cap) {
        if (ce != cudaSuccess) snprintf(err, cap, "svo_dedupe_static: %s: %s", what, cudaGetErrorString(ce));
        else snprintf(err, cap, "svo_dedupe_static: %s", what);
    }
    if (ce != cudaSuccess) cudaGetLastError();
    return code;
}

// can a kernel on the current device dereference p? (device / managed memory, or pinned host
// memory mapped at the same address)
bool ds_device_can_use(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return true;
    return a.type == cudaMemoryTypeHost && a.devicePointer == p;
}

}  // namespace

extern "C" int svo_dedupe_static(int32_t device, const svo_float2* det, int32_t n_det,
                                 const svo_float2* tracked, int32_t n_tracked, float min_distance,
                                 svo_float2* out_xy, int32_t* out_keep_idx, int32_t* out_n,
                                 void* stream, int32_t wait, char* err, size_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (n_det < 0 || n_tracked < 0 || !out_n || (n_det > 0 && !det) || (n_tracked > 0 && !tracked))
        return ds_fail(err, err_cap, SVO_ERR_INVALID, "bad arguments (negative count or null pointer)");
    if (n_det > DS_MAX_DET) {
        char msg[160];
        snprintf(msg, sizeof msg, "n_det = %d exceeds the static scratch (SVO_DEDUPE_STATIC_MAX_DET = %d detections per call)",
                 (int)n_det, DS_MAX_DET);
        return ds_fail(err, err_cap, SVO_ERR_CAPACITY, msg);
    }
    cudaError_t ce;
    if (device < 0) {
        int cur = 0;
        if ((ce = cudaGetDevice(&cur)) != cudaSuccess) return ds_fail(err, err_cap, SVO_ERR_CUDA, "cudaGetDevice", ce);
        device = cur;
    } else if ((ce = cudaSetDevice(device)) != cudaSuccess) {
        return ds_fail(err, err_cap, SVO_ERR_CUDA, "cudaSetDevice", ce);
    }
    if (device >= DS_MAX_DEVICES) return ds_fail(err, err_cap, SVO_ERR_INVALID, "device index out of range");
    if ((n_det > 0 && !ds_device_can_use(det)) || (n_tracked > 0 && !ds_device_can_use(tracked)) ||
        !ds_device_can_use(out_n) || (out_xy && !ds_device_can_use(out_xy)) ||
        (out_keep_idx && !ds_device_can_use(out_keep_idx)))
        return ds_fail(err, err_cap, SVO_ERR_INVALID,
                       "every array must be device memory or pinned host memory (svo_device_alloc / svo_host_alloc): this call stages nothing");
    int* scratch = nullptr;
    if ((ce = ds_scratch_of(device, &scratch)) != cudaSuccess)
        return ds_fail(err, err_cap, SVO_ERR_CUDA, "cudaGetSymbolAddress(g_dedupe_scratch)", ce);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the ticket must start at zero whatever an earlier (possibly aborted) call left behind
    if ((ce = cudaMemsetAsync(scratch, 0, DS_HEADER * sizeof(int), st)) != cudaSuccess)
        return ds_fail(err, err_cap, SVO_ERR_CUDA, "cudaMemsetAsync", ce);
    int blocks = (n_det + DS_THREADS - 1) / DS_THREADS;
    blocks = blocks < 1 ? 1 : (blocks > DS_MAX_BLOCKS ? DS_MAX_BLOCKS : blocks);
    dedupe_static_kernel<<<blocks, DS_THREADS, 0, st>>>(
        reinterpret_cast<const float2*>(det), n_det, reinterpret_cast<const float2*>(tracked),
        n_tracked, min_distance, reinterpret_cast<float2*>(out_xy), out_keep_idx, out_n, scratch);
    if ((ce = cudaGetLastError()) != cudaSuccess) return ds_fail(err, err_cap, SVO_ERR_CUDA, "launch", ce);
    if (wait && (ce = cudaStreamSynchronize(st)) != cudaSuccess)
        return ds_fail(err, err_cap, SVO_ERR_CUDA, "cudaStreamSynchronize", ce);
    return SVO_OK;
}
