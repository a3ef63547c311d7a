// Driver for tests/test_host_sanitizers.py: the pure-host parts of libsvo_host.so (KITTI PNG
// source, multi-threaded decode-ahead ring, image messages + handle_images queue) exercised
// under ThreadSanitizer and AddressSanitizer + UBSan. argv[1] = a KITTI-shaped tree, sequence 00.
#include <cstdio>
#include <cstring>
#include <vector>
#include <stdexcept>
#include "stereo_vo/kitti_source.hpp"
#include "stereo_vo/image_msg.hpp"
int main(int argc, char** argv) {
    const std::string root = argv[1];
    svo::KittiSource src(root, "00");
    const int w = src.width(), h = src.height();
    std::vector<uint8_t> l((size_t)w * h), r((size_t)w * h);
    for (int threads : {1, 3, 8}) {
        for (int depth : {3, 6, 11}) {
            std::vector<uint8_t> slab((size_t)depth * 2 * w * h);
            svo::KittiPrefetcher pf(src, slab.data(), depth, w, 0, threads);
            size_t i; const uint8_t *pl, *pr; size_t n = 0; unsigned long sum = 0;
            const uint8_t* prev = nullptr;
            while (pf.next(i, pl, pr)) {
                if (i != n) { std::printf("order!\n"); return 1; }
                src.read(i, l.data(), r.data(), w);
                if (std::memcmp(pl, l.data(), l.size()) || std::memcmp(pr, r.data(), r.size())) { std::printf("bytes!\n"); return 1; }
                if (prev) sum += prev[17];
                prev = pl;
                ++n;
            }
            std::printf("threads %d depth %d frames %zu ok (%lu)\n", threads, depth, n, sum);
        }
    }
    {   // early destruction
        std::vector<uint8_t> slab((size_t)6 * 2 * w * h);
        svo::KittiPrefetcher pf(src, slab.data(), 6, w, 0, 4);
        size_t i; const uint8_t *pl, *pr;
        pf.next(i, pl, pr);
    }
    // image messages
    auto q = std::make_shared<std::queue<StereoPair>>();
    handle_images on_pair(q, 0.05);
    std::vector<uint8_t> rgb((size_t)w * h * 3, 77);
    for (int k = 0; k < 5; ++k) {
        svo::ImageMsg a{0.03 * k, (uint32_t)h, (uint32_t)w, "mono8", 0, (uint32_t)w, l.data(), l.size()};
        svo::ImageMsg b{0.03 * k, (uint32_t)h, (uint32_t)w, "rgb8", 0, (uint32_t)(3 * w), rgb.data(), rgb.size()};
        on_pair(a, b);
    }
    size_t n = svo::drain_queue(*q, [](const StereoPair& p) { if (p.left.cols <= 0) throw std::runtime_error("x"); });
    std::printf("queued+drained %zu\n", n);
    return 0;
}
