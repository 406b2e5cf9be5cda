// image_io.h -- minimal image file reader of the CLI driver: what `cv::imread(path)` (IMREAD_COLOR, the default)
// gives the reference at src/main.cpp:37-38, for the formats this box can decode without OpenCV:
//   PNG  8-bit gray / RGB / palette / gray+alpha / RGBA, non-interlaced (inflate via zlib)
//   PNM  binary P5 (gray) / P6 (RGB), maxval 255
// The result is always a 3-channel BGR cv::Mat (gray replicated, alpha dropped), like IMREAD_COLOR.
// JPEG needs libjpeg, which is not installed here: read_image() says so.
#ifndef SVO_CLI_IMAGE_IO_H_
#define SVO_CLI_IMAGE_IO_H_

#include <zlib.h>

#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "vo.h"

namespace svo_cli {

inline bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    const size_t got = n > 0 ? std::fread(out.data(), 1, (size_t)n, f) : 0;
    std::fclose(f);
    return got == out.size();
}

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// PNG filter type 4
inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

inline bool decode_png(const std::vector<uint8_t>& file, cv::Mat& bgr, std::string& err) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    if (file.size() < 8 || std::memcmp(file.data(), sig, 8) != 0) { err = "not a PNG"; return false; }
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte;
    size_t pos = 8;
    while (pos + 12 <= file.size()) {
        const uint32_t len = be32(&file[pos]);
        const char* type = reinterpret_cast<const char*>(&file[pos + 4]);
        if (pos + 12 + len > file.size()) { err = "truncated PNG chunk"; return false; }
        const uint8_t* d = &file[pos + 8];
        if (!std::memcmp(type, "IHDR", 4) && len >= 13) {
            w = be32(d); h = be32(d + 4); depth = d[8]; ctype = d[9]; interlace = d[12];
        } else if (!std::memcmp(type, "PLTE", 4)) {
            plte.assign(d, d + len);
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), d, d + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    if (!w || !h || w > 16384 || h > 16384) { err = "bad PNG size"; return false; }
    if (depth != 8 || interlace != 0) { err = "only 8-bit non-interlaced PNGs are supported"; return false; }
    int spp;   // samples per pixel
    switch (ctype) {
        case 0: spp = 1; break;
        case 2: spp = 3; break;
        case 3: spp = 1; break;
        case 4: spp = 2; break;
        case 6: spp = 4; break;
        default: err = "bad PNG colour type"; return false;
    }
    const size_t rowb = (size_t)w * spp;
    std::vector<uint8_t> raw((rowb + 1) * h);
    uLongf rawlen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rawlen, idat.data(), (uLong)idat.size()) != Z_OK || rawlen != raw.size()) {
        err = "PNG inflate failed";
        return false;
    }
    std::vector<uint8_t> prev(rowb, 0), cur(rowb);
    bgr = cv::Mat((int)h, (int)w, CV_8UC3);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t* src = &raw[(rowb + 1) * y];
        const int ft = src[0];
        ++src;
        for (size_t i = 0; i < rowb; ++i) {
            const int a = i >= (size_t)spp ? cur[i - spp] : 0, b = prev[i], c = i >= (size_t)spp ? prev[i - spp] : 0;
            int v = src[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: err = "bad PNG filter"; return false;
            }
            cur[i] = (uint8_t)v;
        }
        uint8_t* dst = bgr.ptr<uint8_t>((int)y);
        for (uint32_t x = 0; x < w; ++x) {
            uint8_t r, g, bl;
            const uint8_t* p = &cur[(size_t)x * spp];
            if (ctype == 2 || ctype == 6) { r = p[0]; g = p[1]; bl = p[2]; }
            else if (ctype == 3) {
                const size_t k = (size_t)p[0] * 3;
                if (k + 2 >= plte.size()) { r = g = bl = 0; }
                else { r = plte[k]; g = plte[k + 1]; bl = plte[k + 2]; }
            } else { r = g = bl = p[0]; }
            dst[3 * x] = bl; dst[3 * x + 1] = g; dst[3 * x + 2] = r;
        }
        prev.swap(cur);
    }
    return true;
}

inline bool decode_pnm(const std::vector<uint8_t>& file, cv::Mat& bgr, std::string& err) {
    if (file.size() < 3 || file[0] != 'P' || (file[1] != '5' && file[1] != '6')) { err = "not a binary PNM"; return false; }
    const int spp = file[1] == '6' ? 3 : 1;
    size_t pos = 2;
    long vals[3];
    for (int k = 0; k < 3; ++k) {
        for (;;) {   // whitespace and comments
            while (pos < file.size() && std::isspace(file[pos])) ++pos;
            if (pos < file.size() && file[pos] == '#') { while (pos < file.size() && file[pos] != '\n') ++pos; }
            else break;
        }
        long v = 0;
        bool any = false;
        while (pos < file.size() && std::isdigit(file[pos])) { v = v * 10 + (file[pos++] - '0'); any = true; }
        if (!any) { err = "bad PNM header"; return false; }
        vals[k] = v;
    }
    ++pos;   // the single whitespace after maxval
    const long w = vals[0], h = vals[1];
    if (vals[2] != 255 || w <= 0 || h <= 0 || w > 16384 || h > 16384 || pos + (size_t)w * h * spp > file.size()) {
        err = "unsupported PNM";
        return false;
    }
    bgr = cv::Mat((int)h, (int)w, CV_8UC3);
    for (long y = 0; y < h; ++y) {
        const uint8_t* p = &file[pos + (size_t)y * w * spp];
        uint8_t* dst = bgr.ptr<uint8_t>((int)y);
        for (long x = 0; x < w; ++x) {
            if (spp == 3) { dst[3 * x] = p[3 * x + 2]; dst[3 * x + 1] = p[3 * x + 1]; dst[3 * x + 2] = p[3 * x]; }
            else dst[3 * x] = dst[3 * x + 1] = dst[3 * x + 2] = p[x];
        }
    }
    return true;
}

// cv::imread(path): empty Mat when the file does not exist (the reference skips such frames, main.cpp:376-378)
inline cv::Mat read_image(const std::string& path, std::string* why = nullptr) {
    std::vector<uint8_t> file;
    cv::Mat out;
    std::string err;
    if (!read_file(path, file)) { if (why) *why = "cannot open " + path; return out; }
    bool ok;
    if (file.size() >= 2 && file[0] == 0xFF && file[1] == 0xD8) { ok = false; err = "JPEG decoding is unavailable on this box (no libjpeg)"; }
    else if (file.size() >= 2 && file[0] == 'P') ok = decode_pnm(file, out, err);
    else ok = decode_png(file, out, err);
    if (!ok) { if (why) *why = path + ": " + err; return cv::Mat(); }
    return out;
}

// cv::cvtColor(BGR2GRAY) for 8-bit as OpenCV 4.x computes it: fixed point with 15 fractional bits, coefficients
// 3735 / 19235 / 9798 (B, G, R), round to nearest (tests/test_cli.py compares with cv2.cvtColor)
inline cv::Mat bgr_to_gray(const cv::Mat& bgr) {
    cv::Mat g(bgr.rows, bgr.cols, CV_8UC1);
    for (int y = 0; y < bgr.rows; ++y) {
        const uint8_t* p = bgr.ptr<uint8_t>(y);
        uint8_t* d = g.ptr<uint8_t>(y);
        for (int x = 0; x < bgr.cols; ++x) d[x] = (uint8_t)((p[3 * x] * 3735 + p[3 * x + 1] * 19235 + p[3 * x + 2] * 9798 + (1 << 14)) >> 15);
    }
    return g;
}

}  // namespace svo_cli
#endif
