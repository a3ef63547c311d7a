// track_filter_compact: the filter + statistics tail of FeatureTracker::track_features
// (reference src/feature_tracker.cpp:38-66).
//   keep i  <=>  status1[i] && (!flow_back || (status2[i] && norm(old[i]-reverse[i]) < 2))
//                && parallax_i <= 200,   parallax_i = |tracked[i] - initial_features[id_i]| (f32)
//   av_parallax  = (sum of kept parallaxes, f32, in index order) / n_attempted     
//   percent_lost = 1.0 - kept / (float)initial_features.size()                     
// The kept features, their ids (id = old_feature_ids[i]) and their initial positions
// are compacted IN ORDER into the other feature buffer of the sequence; one block per sequence,
// ballot/prefix compaction, the f32 sum is done sequentially by one thread so the keyframe
// decision that depends on it (src/image_processor.cpp:63) is reproducible.
#include "svo_common.cuh"

namespace {

constexpr int TF_THREADS = 256;
constexpr unsigned FULL = 0xffffffffu;

__global__ void __launch_bounds__(TF_THREADS)
track_filter_kernel(float2* __restrict__ feat, unsigned long long* __restrict__ ids,
                    float2* __restrict__ initial, size_t feat_stride,
                    const float2* __restrict__ tracked, const unsigned char* __restrict__ status1,
                    const float2* __restrict__ reverse, const unsigned char* __restrict__ status2,
                    int out_stride, int flow_back, double fb_max, float max_parallax,
                    int* __restrict__ kept_idx, float* __restrict__ kept_parallax,
                    SvoTrackStats* __restrict__ stats, int max_features,
                    const int* __restrict__ gate, int gate_min,
                    const SvoSeqCall* __restrict__ call) {
    __shared__ int s_w[TF_THREADS / 32];
    const SvoSeqCall cd = call[blockIdx.x];
    const int n = cd.n_feat;
    if ((cd.flags & 1) == 0 || (gate && gate[cd.seq] < gate_min)) {
        // tracker not initialised, or < 4 corners (image_processor.cpp:23-25): not advanced
        if (threadIdx.x == 0) {
            SvoTrackStats s = {0.f, 0.f, 0, -1};
            stats[cd.seq] = s;
        }
        return;
    }
    const size_t ib = (size_t)cd.seq * feat_stride + (size_t)cd.feat_buf * max_features;
    const size_t ob = (size_t)cd.seq * feat_stride + (size_t)(1 - cd.feat_buf) * max_features;
    const size_t tb = (size_t)cd.seq * out_stride;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    int written = 0;
    for (int i0 = 0; i0 < n; i0 += TF_THREADS) {
        const int i = i0 + threadIdx.x;
        bool keep = false;
        float parallax = 0.f;
        float2 t = make_float2(0.f, 0.f), ini = t;
        unsigned long long id = 0;
        if (i < n) {
            t = tracked[tb + i];
            ini = initial[ib + i];
            id = ids[ib + i];
            keep = status1[tb + i] != 0;
            if (keep && flow_back) {
                float2 o = feat[ib + i], r = reverse[tb + i];
                float dx = __fsub_rn(o.x, r.x), dy = __fsub_rn(o.y, r.y);   // Point2f difference
                double nrm = sqrt((double)dx * (double)dx + (double)dy * (double)dy);  // cv::norm
                keep = (status2[tb + i] != 0) && (nrm < fb_max);
            }
            if (keep) {
                float dx = __fsub_rn(t.x, ini.x), dy = __fsub_rn(t.y, ini.y);
                parallax = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
                if (parallax > max_parallax) keep = false;
            }
        }
        unsigned bal = __ballot_sync(FULL, keep);
        if (lane == 0) s_w[wrp] = __popc(bal);
        __syncthreads();
        int off = 0, total = 0;
        for (int w = 0; w < TF_THREADS / 32; ++w) {
            int c = s_w[w];
            if (w < wrp) off += c;
            total += c;
        }
        __syncthreads();
        if (keep) {
            int slot = written + off + __popc(bal & ((1u << lane) - 1u));
            feat[ob + slot] = t;
            ids[ob + slot] = id;
            initial[ob + slot] = ini;
            kept_idx[tb + slot] = i;
            kept_parallax[tb + slot] = parallax;
        }
        written += total;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float av = 0.f;
        for (int k = 0; k < written; ++k) av = __fadd_rn(av, kept_parallax[tb + k]);
        SvoTrackStats s;
        s.av_parallax = __fdiv_rn(av, (float)n);                       // n == 0 -> NaN, as the reference
        float ratio = __fdiv_rn((float)written, (float)cd.n_initial);
        s.percent_lost = (float)(1.0 - (double)ratio);
        s.n_kept = written;
        s.n_attempted = n;
        stats[cd.seq] = s;
    }
}

}  // namespace

cudaError_t svo_launch_track_filter(float2* feat, unsigned long long* ids,
                                    float2* initial, size_t feat_stride,
                                    const float2* tracked, const unsigned char* status1,
                                    const float2* reverse, const unsigned char* status2,
                                    int out_stride, int flow_back, float fb_max, float max_parallax,
                                    int* kept_idx, float* kept_parallax, SvoTrackStats* stats,
                                    int max_features, const int* gate, int gate_min,
                                    const SvoSeqCall* d_call, int n_call, cudaStream_t st) {
    track_filter_kernel<<<n_call, TF_THREADS, 0, st>>>(
        feat, ids, initial, feat_stride, tracked, status1, reverse, status2, out_stride,
        flow_back, (double)fb_max, max_parallax, kept_idx, kept_parallax, stats, max_features,
        gate, gate_min, d_call);
    return cudaGetLastError();
}
