// KITTI odometry frame source (see include/stereo_vo/kitti_source.hpp): pose file parser, a
// small PNG reader on zlib, and a decode-ahead thread. Host I/O only; no pixel arithmetic of the
// front end lives here.
#include "stereo_vo/kitti_source.hpp"

#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace svo {

namespace {

struct PngInfo {
    int width = 0, height = 0, bit_depth = 0, color_type = 0, interlace = 0;
};

uint32_t be32(const unsigned char* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3];
}

[[noreturn]] void bad(const std::string& path, const char* what) {
    throw std::runtime_error("PNG " + path + ": " + what);
}

// whole file -> IHDR fields + concatenated IDAT payload
PngInfo parse_png(const std::string& path, std::vector<unsigned char>* idat) {
    std::ifstream f(path, std::ios::binary);
    if (!f) bad(path, "cannot open");
    std::vector<unsigned char> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (buf.size() < 8 + 25 || std::memcmp(buf.data(), sig, 8) != 0) bad(path, "not a PNG file");
    PngInfo info;
    size_t pos = 8;
    bool have_ihdr = false, done = false;
    while (!done && pos + 12 <= buf.size()) {
        const uint32_t len = be32(&buf[pos]);
        const unsigned char* type = &buf[pos + 4];
        if (pos + 12 + (size_t)len > buf.size()) bad(path, "truncated chunk");
        const unsigned char* data = &buf[pos + 8];
        if (std::memcmp(type, "IHDR", 4) == 0) {
            if (len != 13) bad(path, "bad IHDR");
            info.width = (int)be32(data);
            info.height = (int)be32(data + 4);
            info.bit_depth = data[8];
            info.color_type = data[9];
            info.interlace = data[12];
            if (info.width <= 0 || info.height <= 0 || info.width > 65536 || info.height > 65536)
                bad(path, "implausible image size in IHDR");
            have_ihdr = true;
            if (!idat) return info;
        } else if (std::memcmp(type, "IDAT", 4) == 0) {
            if (idat) idat->insert(idat->end(), data, data + len);
        } else if (std::memcmp(type, "IEND", 4) == 0) {
            done = true;
        }
        pos += 12 + (size_t)len;
    }
    if (!have_ihdr) bad(path, "no IHDR");
    return info;
}

int channels_of(int color_type) {
    switch (color_type) {
        case 0: return 1;
        case 2: return 3;
        case 4: return 2;
        case 6: return 4;
        default: return 0;
    }
}

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

void read_png_size(const std::string& path, int& width, int& height) {
    const PngInfo info = parse_png(path, nullptr);
    width = info.width;
    height = info.height;
}

void read_png_gray(const std::string& path, uint8_t* out, int pitch, int width, int height) {
    std::vector<unsigned char> idat;
    const PngInfo info = parse_png(path, &idat);
    if (info.width != width || info.height != height) bad(path, "image size differs from the sequence's");
    if (info.bit_depth != 8) bad(path, "only 8-bit PNGs are supported");
    if (info.interlace != 0) bad(path, "interlaced PNGs are not supported");
    const int ch = channels_of(info.color_type);
    if (ch == 0) bad(path, "palette PNGs are not supported");
    if (pitch < width) bad(path, "pitch < width");
    const size_t stride = (size_t)width * ch;
    std::vector<unsigned char> raw((stride + 1) * (size_t)height);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size())
        bad(path, "zlib inflate failed");
    std::vector<unsigned char> prev(stride, 0), cur(stride);
    for (int y = 0; y < height; ++y) {
        const unsigned char* line = &raw[(stride + 1) * (size_t)y];
        const int filter = line[0];
        const unsigned char* s = line + 1;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= (size_t)ch ? cur[i - ch] : 0, b = prev[i], c = i >= (size_t)ch ? prev[i - ch] : 0;
            int v = s[i];
            switch (filter) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: bad(path, "unknown filter type");
            }
            cur[i] = (unsigned char)v;
        }
        uint8_t* o = out + (size_t)y * pitch;
        if (ch == 1) {
            std::memcpy(o, cur.data(), (size_t)width);
        } else if (ch == 2) {
            for (int x = 0; x < width; ++x) o[x] = cur[2 * (size_t)x];
        } else {
            // cv::imread(..., GRAYSCALE) lets libpng convert: png_set_rgb_to_gray(0.299, 0.587)
            // = 15-bit coefficients 9797 / 19234 / 3737, truncated (checked against cv2.imread)
            for (int x = 0; x < width; ++x) {
                const unsigned char* px = &cur[(size_t)x * ch];
                o[x] = (uint8_t)((px[0] * 9797 + px[1] * 19234 + px[2] * 3737) >> 15);
            }
        }
        prev.swap(cur);
    }
}

KittiSource::KittiSource(const std::string& data_path, const std::string& sequence)
    : root_(data_path), seq_(sequence) {
    // src/kitti_node.cpp:37: data_path + "data_odometry_poses/dataset/poses/" + sequence + ".txt"
    const std::string pose_file = root_ + "data_odometry_poses/dataset/poses/" + seq_ + ".txt";
    std::ifstream gt(pose_file);
    if (!gt) throw std::runtime_error("KittiSource: cannot open " + pose_file);
    KittiPose p;
    // :47-50: the loop runs while twelve numbers can be read; a short last line ends it
    while (gt >> p.R[0] >> p.R[1] >> p.R[2] >> p.t[0] >> p.R[3] >> p.R[4] >> p.R[5] >> p.t[1] >>
           p.R[6] >> p.R[7] >> p.R[8] >> p.t[2])
        poses_.push_back(p);
    if (poses_.empty()) throw std::runtime_error("KittiSource: no poses in " + pose_file);
    read_png_size(left_path(0), width_, height_);
}

std::string KittiSource::image_path(int cam, size_t i) const {
    // :56-68: six digits, zero padded (the reference's prefix ladder breaks at 100000 frames;
    // no KITTI sequence has that many)
    char name[32];
    std::snprintf(name, sizeof name, "%06zu.png", i);
    return root_ + seq_ + (cam == 0 ? "/image_0/" : "/image_1/") + name;
}

void KittiSource::read(size_t i, uint8_t* left, uint8_t* right, int pitch) const {
    if (i >= poses_.size()) throw std::out_of_range("KittiSource::read: frame index past the pose file");
    read_png_gray(left_path(i), left, pitch, width_, height_);
    read_png_gray(right_path(i), right, pitch, width_, height_);
}

KittiPrefetcher::KittiPrefetcher(const KittiSource& src, uint8_t* slab, int depth, int pitch, size_t first)
    : src_(src), slab_(slab), depth_(depth), pitch_(pitch), produced_(first), consumed_(first),
      end_(src.size()) {
    if (!slab || depth < 3 || pitch < src.width()) throw std::invalid_argument("KittiPrefetcher: bad arguments");
    worker_ = std::thread(&KittiPrefetcher::run, this);
}

KittiPrefetcher::~KittiPrefetcher() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_.notify_all();
    if (worker_.joinable()) worker_.join();
}

void KittiPrefetcher::run() {
    const size_t slot_bytes = (size_t)2 * pitch_ * src_.height();
    for (;;) {
        size_t i;
        {
            std::unique_lock<std::mutex> lk(mu_);
            // the two frames handed out last stay untouched (two steps may be in flight)
            cv_.wait(lk, [&] { return stop_ || (produced_ < end_ && produced_ - consumed_ < (size_t)depth_ - 2); });
            if (stop_) return;
            i = produced_;
        }
        uint8_t* slot = slab_ + (i % (size_t)depth_) * slot_bytes;
        try {
            src_.read(i, slot, slot + (size_t)pitch_ * src_.height(), pitch_);
        } catch (const std::exception& ex) {
            std::lock_guard<std::mutex> lk(mu_);
            error_ = ex.what();
            stop_ = true;
            cv_.notify_all();
            return;
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            produced_ = i + 1;
        }
        cv_.notify_all();
    }
}

bool KittiPrefetcher::next(size_t& index, const uint8_t*& left, const uint8_t*& right) {
    std::unique_lock<std::mutex> lk(mu_);
    if (consumed_ >= end_) return false;
    cv_.wait(lk, [&] { return produced_ > consumed_ || !error_.empty(); });
    if (produced_ <= consumed_) throw std::runtime_error(error_);
    index = consumed_;
    const size_t slot_bytes = (size_t)2 * pitch_ * src_.height();
    left = slab_ + (index % (size_t)depth_) * slot_bytes;
    right = left + (size_t)pitch_ * src_.height();
    ++consumed_;
    lk.unlock();
    cv_.notify_all();
    return true;
}

}  // namespace svo
