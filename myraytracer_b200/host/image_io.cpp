// image_io.cpp — headless image output of the host mirror.
//
// The reference turns the RGBA8 frame into a CGImage and writes PNG files through ImageIO
// (Helpers/Image.swift: makeCGImage / writePNG; BatchRender.swift encodes frames with AVFoundation) — Apple
// frameworks that do not exist here.  The frame is a plain buffer in this module, written as
//   * PNG  — 8-bit RGBA, one IDAT of stored (uncompressed) deflate blocks: no zlib dependency, any viewer reads it
//   * PPM  — binary P6 (RGB, alpha dropped)
// chosen by the file extension.  Row 0 is the top row in both formats, as in the frame (RayTracer.swift:186-195).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rtcuda_host.h"

namespace {

uint32_t crcTable[256];
bool crcReady = false;
uint32_t crc32(uint32_t crc, const uint8_t* p, size_t n) {
    if (!crcReady) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            crcTable[i] = c;
        }
        crcReady = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = crcTable[(crc ^ p[i]) & 0xffu] ^ (crc >> 8);
    return ~crc;
}

void be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }

bool writeChunk(FILE* f, const char type[4], const std::vector<uint8_t>& data) {
    std::vector<uint8_t> head;
    be32(head, (uint32_t)data.size());
    uint8_t t[4];
    std::memcpy(t, type, 4);
    uint32_t crc = crc32(0, t, 4);
    if (!data.empty()) crc = crc32(crc, data.data(), data.size());
    std::vector<uint8_t> tail;
    be32(tail, crc);
    return std::fwrite(head.data(), 1, 4, f) == 4 && std::fwrite(t, 1, 4, f) == 4 &&
           (data.empty() || std::fwrite(data.data(), 1, data.size(), f) == data.size()) && std::fwrite(tail.data(), 1, 4, f) == 4;
}

bool writePng(FILE* f, const uint8_t* rgba, int w, int h) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (std::fwrite(sig, 1, 8, f) != 8) return false;
    std::vector<uint8_t> ihdr;
    be32(ihdr, (uint32_t)w); be32(ihdr, (uint32_t)h);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8-bit RGBA, no interlace
    if (!writeChunk(f, "IHDR", ihdr)) return false;
    // raw scanlines: filter byte 0 + row
    const size_t stride = (size_t)w * 4 + 1;
    std::vector<uint8_t> raw(stride * (size_t)h);
    for (int y = 0; y < h; y++) {
        raw[y * stride] = 0;
        std::memcpy(&raw[y * stride + 1], rgba + (size_t)y * w * 4, (size_t)w * 4);
    }
    // zlib stream of stored blocks (<= 65535 bytes each) + Adler-32
    std::vector<uint8_t> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    size_t pos = 0;
    do {
        const size_t n = raw.size() - pos < 65535 ? raw.size() - pos : 65535;
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back(n & 0xff); z.push_back(n >> 8); z.push_back(~n & 0xff); z.push_back((~n >> 8) & 0xff);
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        for (size_t i = 0; i < n; i++) { a = (a + raw[pos + i]) % 65521u; b = (b + a) % 65521u; }
        pos += n;
    } while (pos < raw.size());
    be32(z, (b << 16) | a);
    return writeChunk(f, "IDAT", z) && writeChunk(f, "IEND", {});
}

bool writePpm(FILE* f, const uint8_t* rgba, int w, int h) {
    if (std::fprintf(f, "P6\n%d %d\n255\n", w, h) < 0) return false;
    std::vector<uint8_t> row((size_t)w * 3);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) std::memcpy(&row[(size_t)x * 3], rgba + ((size_t)y * w + x) * 4, 3);
        if (std::fwrite(row.data(), 1, row.size(), f) != row.size()) return false;
    }
    return true;
}

}  // namespace

extern "C" int32_t rth_image_write(const char* path, const uint8_t* rgba, int32_t width, int32_t height) {
    if (!path || !rgba || width < 1 || height < 1) return RT_ERR_INVALID;
    const std::string p(path);
    const size_t dot = p.rfind('.');
    std::string ext = dot == std::string::npos ? "" : p.substr(dot + 1);
    for (char& c : ext) c = (char)std::tolower((unsigned char)c);
    if (ext != "png" && ext != "ppm") return RT_ERR_UNSUPPORTED;
    FILE* f = std::fopen(path, "wb");
    if (!f) return RT_ERR_INVALID;
    const bool ok = ext == "png" ? writePng(f, rgba, width, height) : writePpm(f, rgba, width, height);
    return (std::fclose(f) == 0 && ok) ? RT_OK : RT_ERR_INVALID;
}
