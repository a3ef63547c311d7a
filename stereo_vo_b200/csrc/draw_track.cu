// draw_track: FeatureTracker::draw_track of the reference (src/feature_tracker.cpp:74-83) -
//   track_drawing = initial_image.clone(); cvtColor(GRAY2RGB);
//   for every tracked feature: cv::arrowedLine(drawing, initial, current, CV_RGB(0, 255, 0), 4)
// - as two kernels: the grey image replicated into three channels, then one block per feature
// that paints the arrow's three thick segments (shaft + two tip strokes, cv::arrowedLine's
// geometry: tips of length 0.1 |shaft| at +-45 degrees, end points rounded to integer pixels,
// thickness 4) into the RGB image. A pixel is painted when its centre lies within thickness / 2
// of a segment; overlapping arrows write the same colour, so the order of blocks does not
// matter. OpenCV's fixed-point polygon + end-cap rasteriser differs from this distance test by
// at most a pixel along the outline (tests compare the painted masks), the picture is the same.
// (the overlay vo_node publishes on /feature_tracking, vo_node.cpp:188.)
#include "svo_common.cuh"

namespace {

__global__ void __launch_bounds__(256)
gray_to_rgb_kernel(const uint8_t* __restrict__ gray, int gpitch, uint8_t* __restrict__ rgb,
                   int rpitch, int w, int h) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w || y >= h) return;
    const uint8_t v = gray[(size_t)y * gpitch + x];
    uint8_t* o = rgb + (size_t)y * rpitch + 3 * x;
    o[0] = v; o[1] = v; o[2] = v;
}

__device__ __forceinline__ float seg_dist2(float px, float py, float ax, float ay, float bx, float by) {
    const float dx = bx - ax, dy = by - ay;
    const float l2 = dx * dx + dy * dy;
    float t = l2 > 0.f ? ((px - ax) * dx + (py - ay) * dy) / l2 : 0.f;
    t = fminf(fmaxf(t, 0.f), 1.f);
    const float qx = ax + t * dx - px, qy = ay + t * dy - py;
    return qx * qx + qy * qy;
}

__global__ void __launch_bounds__(128)
draw_arrows_kernel(const float2* __restrict__ from, const float2* __restrict__ to, int n,
                   uint8_t* __restrict__ rgb, int rpitch, int w, int h, float half_thickness,
                   float tip_length) {
    const int i = blockIdx.x;
    if (i >= n) return;
    // cv::arrowedLine takes cv::Point (int): Point2f -> Point rounds half to even (cvRound)
    const float ax = rintf(from[i].x), ay = rintf(from[i].y);
    const float bx = rintf(to[i].x), by = rintf(to[i].y);
    const float len = sqrtf((ax - bx) * (ax - bx) + (ay - by) * (ay - by));
    const float tip = len * tip_length;
    const float ang = atan2f(ay - by, ax - bx);
    const float k = 0.78539816339744831f;             // CV_PI / 4
    const float t1x = rintf(bx + tip * cosf(ang + k)), t1y = rintf(by + tip * sinf(ang + k));
    const float t2x = rintf(bx + tip * cosf(ang - k)), t2y = rintf(by + tip * sinf(ang - k));
    const float r = half_thickness, r2 = r * r;
    const int x0 = max(0, (int)floorf(fminf(fminf(ax, bx), fminf(t1x, t2x)) - r - 1.f));
    const int x1 = min(w - 1, (int)ceilf(fmaxf(fmaxf(ax, bx), fmaxf(t1x, t2x)) + r + 1.f));
    const int y0 = max(0, (int)floorf(fminf(fminf(ay, by), fminf(t1y, t2y)) - r - 1.f));
    const int y1 = min(h - 1, (int)ceilf(fmaxf(fmaxf(ay, by), fmaxf(t1y, t2y)) + r + 1.f));
    const int bw = x1 - x0 + 1, bh = y1 - y0 + 1;
    if (bw <= 0 || bh <= 0) return;
    for (int p = threadIdx.x; p < bw * bh; p += blockDim.x) {
        const int x = x0 + p % bw, y = y0 + p / bw;
        const float fx = (float)x, fy = (float)y;
        const bool on = seg_dist2(fx, fy, ax, ay, bx, by) <= r2 ||
                        seg_dist2(fx, fy, t1x, t1y, bx, by) <= r2 ||
                        seg_dist2(fx, fy, t2x, t2y, bx, by) <= r2;
        if (on) {
            uint8_t* o = rgb + (size_t)y * rpitch + 3 * x;
            o[0] = 0; o[1] = 255; o[2] = 0;              // CV_RGB(0, 255, 0) in an RGB image
        }
    }
}

}  // namespace

cudaError_t svo_launch_draw_track(const uint8_t* gray, int gpitch, const float2* from,
                                  const float2* to, int n, uint8_t* rgb, int rpitch, int w, int h,
                                  int thickness, cudaStream_t st) {
    dim3 grid((w + 255) / 256, h);
    gray_to_rgb_kernel<<<grid, 256, 0, st>>>(gray, gpitch, rgb, rpitch, w, h);
    if (n > 0)
        draw_arrows_kernel<<<n, 128, 0, st>>>(from, to, n, rgb, rpitch, w, h, 0.5f * (float)thickness, 0.1f);
    return cudaGetLastError();
}
