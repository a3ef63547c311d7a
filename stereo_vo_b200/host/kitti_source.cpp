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

// whole file in memory
std::vector<unsigned char> slurp(const std::string& path) {
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) bad(path, "cannot open");
    std::vector<unsigned char> buf;
    if (std::fseek(f, 0, SEEK_END) == 0) {
        const long n = std::ftell(f);
        if (n > 0) buf.resize((size_t)n);
        std::rewind(f);
    }
    size_t got = buf.empty() ? 0 : std::fread(buf.data(), 1, buf.size(), f);
    if (buf.empty()) {                                   // not seekable: read in pieces
        unsigned char piece[1 << 16];
        size_t k;
        while ((k = std::fread(piece, 1, sizeof piece, f)) > 0) buf.insert(buf.end(), piece, piece + k);
        got = buf.size();
    }
    std::fclose(f);
    if (got != buf.size()) bad(path, "short read");
    return buf;
}

struct Span {
    const unsigned char* p;
    size_t n;
};

// file bytes -> IHDR fields + the IDAT payload pieces in file order (pointers into buf)
PngInfo parse_png(const std::string& path, const std::vector<unsigned char>& buf, std::vector<Span>* idat) {
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
            if (idat && len > 0) idat->push_back(Span{data, len});
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

// Undo one row's PNG filter: s = the filtered bytes, d = where the row goes, up = the row above
// (already unfiltered; nullptr for the first row = all zeros), n bytes, bpp bytes per pixel.
// One loop per filter type; the first pixel (no left neighbour) is peeled off.
bool unfilter_row(int filter, const unsigned char* s, unsigned char* d, const unsigned char* up,
                  size_t n, size_t bpp) {
    const size_t head = bpp < n ? bpp : n;
    switch (filter) {
        case 0:
            std::memcpy(d, s, n);
            return true;
        case 1:
            std::memcpy(d, s, head);
            for (size_t i = head; i < n; ++i) d[i] = (unsigned char)(s[i] + d[i - bpp]);
            return true;
        case 2:
            if (!up) { std::memcpy(d, s, n); return true; }
            for (size_t i = 0; i < n; ++i) d[i] = (unsigned char)(s[i] + up[i]);
            return true;
        case 3:
            if (!up) {
                std::memcpy(d, s, head);
                for (size_t i = head; i < n; ++i) d[i] = (unsigned char)(s[i] + (d[i - bpp] >> 1));
                return true;
            }
            for (size_t i = 0; i < head; ++i) d[i] = (unsigned char)(s[i] + (up[i] >> 1));
            for (size_t i = head; i < n; ++i) d[i] = (unsigned char)(s[i] + ((d[i - bpp] + up[i]) >> 1));
            return true;
        case 4:
            if (!up) {                                   // paeth(a, 0, 0) = a
                std::memcpy(d, s, head);
                for (size_t i = head; i < n; ++i) d[i] = (unsigned char)(s[i] + d[i - bpp]);
                return true;
            }
            for (size_t i = 0; i < head; ++i) d[i] = (unsigned char)(s[i] + up[i]);   // paeth(0, b, 0) = b
            for (size_t i = head; i < n; ++i)
                d[i] = (unsigned char)(s[i] + paeth(d[i - bpp], up[i], up[i - bpp]));
            return true;
        default:
            return false;
    }
}

}  // namespace

void read_png_size(const std::string& path, int& width, int& height) {
    const std::vector<unsigned char> buf = slurp(path);
    const PngInfo info = parse_png(path, buf, nullptr);
    width = info.width;
    height = info.height;
}

void read_png_gray(const std::string& path, uint8_t* out, int pitch, int width, int height) {
    const std::vector<unsigned char> buf = slurp(path);
    std::vector<Span> idat;
    const PngInfo info = parse_png(path, buf, &idat);
    if (info.width != width || info.height != height) bad(path, "image size differs from the sequence's");
    if (info.bit_depth != 8) bad(path, "only 8-bit PNGs are supported");
    if (info.interlace != 0) bad(path, "interlaced PNGs are not supported");
    const int ch = channels_of(info.color_type);
    if (ch == 0) bad(path, "palette PNGs are not supported");
    if (pitch < width) bad(path, "pitch < width");
    const size_t stride = (size_t)width * ch;
    // inflate the IDAT pieces where they lie (no concatenated copy) into filter byte + row, ...
    if ((stride + 1) * (size_t)height > (size_t)0x7fffffff) bad(path, "image too large");
    std::vector<unsigned char> raw((stride + 1) * (size_t)height);
    z_stream zs;
    std::memset(&zs, 0, sizeof zs);
    if (inflateInit(&zs) != Z_OK) bad(path, "zlib inflate failed");
    zs.next_out = raw.data();
    zs.avail_out = (uInt)raw.size();
    int zrc = Z_OK;
    for (size_t k = 0; k < idat.size() && zrc == Z_OK; ++k) {
        zs.next_in = const_cast<Bytef*>(idat[k].p);
        zs.avail_in = (uInt)idat[k].n;
        zrc = inflate(&zs, Z_NO_FLUSH);
        if (zrc == Z_OK && zs.avail_in != 0) zrc = Z_BUF_ERROR;      // more pixels than the header says
    }
    const bool whole = zrc == Z_STREAM_END && zs.total_out == raw.size();
    inflateEnd(&zs);
    if (!whole) bad(path, "zlib inflate failed");
    // ... then undo the row filters: grey rows straight into the caller's image (the row above
    // is the caller's previous row), colour rows in a pair of line buffers and weighted
    std::vector<unsigned char> line[2];
    if (ch > 1) { line[0].resize(stride); line[1].resize(stride); }
    for (int y = 0; y < height; ++y) {
        const unsigned char* src = &raw[(stride + 1) * (size_t)y];
        uint8_t* o = out + (size_t)y * pitch;
        if (ch == 1) {
            if (!unfilter_row(src[0], src + 1, o, y ? o - pitch : nullptr, stride, 1))
                bad(path, "unknown filter type");
            continue;
        }
        unsigned char* cur = line[y & 1].data();
        if (!unfilter_row(src[0], src + 1, cur, y ? line[(y & 1) ^ 1].data() : nullptr, stride, (size_t)ch))
            bad(path, "unknown filter type");
        if (ch == 2) {
            for (int x = 0; x < width; ++x) o[x] = cur[2 * (size_t)x];
        } else {
            // cv::imread(..., GRAYSCALE) lets libpng convert: png_set_rgb_to_gray(0.299, 0.587)
            // = 15-bit coefficients 9797 / 19234 / 3737, truncated (checked against cv2.imread)
            for (int x = 0; x < width; ++x) {
                const unsigned char* px = &cur[(size_t)x * ch];
                o[x] = (uint8_t)((px[0] * 9797 + px[1] * 19234 + px[2] * 3737) >> 15);
            }
        }
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

KittiPrefetcher::KittiPrefetcher(const KittiSource& src, uint8_t* slab, int depth, int pitch, size_t first,
                                 int threads)
    : src_(src), slab_(slab), depth_(depth), pitch_(pitch), claimed_(first), consumed_(first),
      end_(src.size()) {
    if (!slab || depth < 3 || pitch < src.width() || threads < 1 || threads > 64)
        throw std::invalid_argument("KittiPrefetcher: bad arguments");
    ready_.assign((size_t)depth, 0);
    for (int t = 0; t < threads; ++t) workers_.emplace_back(&KittiPrefetcher::run, this);
}

KittiPrefetcher::~KittiPrefetcher() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_.notify_all();
    for (std::thread& w : workers_)
        if (w.joinable()) w.join();
}

void KittiPrefetcher::run() {
    const size_t slot_bytes = (size_t)2 * pitch_ * src_.height();
    for (;;) {
        size_t i;
        {
            std::unique_lock<std::mutex> lk(mu_);
            // a worker takes the next frame nobody has taken; the two frames handed out last
            // stay untouched (two steps may be in flight), so frame i may be written once
            // i - consumed_ < depth - 2
            cv_.wait(lk, [&] { return stop_ || (claimed_ < end_ && claimed_ - consumed_ < (size_t)depth_ - 2); });
            if (stop_) return;
            i = claimed_++;
        }
        uint8_t* slot = slab_ + (i % (size_t)depth_) * slot_bytes;
        try {
            src_.read(i, slot, slot + (size_t)pitch_ * src_.height(), pitch_);
        } catch (const std::exception& ex) {
            // frames before this one were taken earlier and will still be delivered; next()
            // throws when it reaches the first frame that failed
            std::lock_guard<std::mutex> lk(mu_);
            if (error_.empty() || i < error_index_) {
                error_ = ex.what();
                error_index_ = i;
            }
            stop_ = true;
            cv_.notify_all();
            return;
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            ready_[i % (size_t)depth_] = i + 1;
        }
        cv_.notify_all();
    }
}

bool KittiPrefetcher::next(size_t& index, const uint8_t*& left, const uint8_t*& right) {
    std::unique_lock<std::mutex> lk(mu_);
    if (consumed_ >= end_) return false;
    const size_t want = consumed_, slot_no = want % (size_t)depth_;
    cv_.wait(lk, [&] { return ready_[slot_no] == want + 1 || (!error_.empty() && want >= error_index_); });
    if (ready_[slot_no] != want + 1) throw std::runtime_error(error_);
    index = want;
    const size_t slot_bytes = (size_t)2 * pitch_ * src_.height();
    left = slab_ + slot_no * slot_bytes;
    right = left + (size_t)pitch_ * src_.height();
    ++consumed_;
    lk.unlock();
    cv_.notify_all();
    return true;
}

}  // namespace svo
