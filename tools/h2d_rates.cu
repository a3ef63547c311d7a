// H2D copy-rate probe for the ingest path: how fast do 16 KITTI-sized images reach the device
// as separate copies, as one strided (2D) copy, and as one contiguous copy?
//   nvcc -O2 -o /tmp/h2d_rates tools/h2d_rates.cu && /tmp/h2d_rates
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)
int main() {
    const size_t img = 1241 * 376, n = 16, far = 12 * 2 * img;
    unsigned char *h, *d;
    CK(cudaHostAlloc(&h, n * far + img, cudaHostAllocMapped | cudaHostAllocPortable));
    CK(cudaMalloc(&d, n * 2 * img));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    auto run = [&](const char* name, auto body) {
        float best = 1e9f;
        for (int rep = 0; rep < 20; ++rep) {
            CK(cudaEventRecord(a, st));
            for (int k = 0; k < 8; ++k) body();
            CK(cudaEventRecord(b, st));
            CK(cudaEventSynchronize(b));
            float ms; CK(cudaEventElapsedTime(&ms, a, b));
            if (ms < best) best = ms;
        }
        printf("%-44s %7.1f us per 16 images  %5.1f GB/s\n", name, best / 8 * 1e3, 8 * n * img / (best * 1e-3) / 1e9);
    };
    run("16 copies of 466 KB, sources 11 MB apart", [&] {
        for (size_t i = 0; i < n; ++i) CK(cudaMemcpyAsync(d + i * 2 * img, h + i * far, img, cudaMemcpyHostToDevice, st)); });
    run("one 2D copy, 16 rows 11 MB apart", [&] {
        CK(cudaMemcpy2DAsync(d, 2 * img, h, far, img, n, cudaMemcpyHostToDevice, st)); });
    run("two 2D copies, 8 rows 11 MB apart", [&] {
        CK(cudaMemcpy2DAsync(d, 2 * img, h, far, img, 8, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpy2DAsync(d + 16 * img, 2 * img, h + 8 * far, far, img, 8, cudaMemcpyHostToDevice, st)); });
    run("one 2D copy, 16 rows 933 KB apart (pairs)", [&] {
        CK(cudaMemcpy2DAsync(d, 2 * img, h, 2 * img, img, n, cudaMemcpyHostToDevice, st)); });
    run("one contiguous copy of 16 images", [&] {
        CK(cudaMemcpyAsync(d, h, n * img, cudaMemcpyHostToDevice, st)); });
    run("one contiguous copy of 16 images, 2D dst", [&] {
        CK(cudaMemcpy2DAsync(d, 2 * img, h, img, img, n, cudaMemcpyHostToDevice, st)); });
    return 0;
}
