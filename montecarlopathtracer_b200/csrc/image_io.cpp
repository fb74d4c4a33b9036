// Image writers for the headless driver (host, product code).
//   PNG : 8-bit RGB of pow(mean radiance, 1/2.2) clamped to [0,1] -- the display value the reference keeps
//         in its accumulation texture (rtx.hlsl:401-403) and dumps to temp.png / step%06d.png
//         (GraphicsRTX.cpp:223-226, CUTracer.cu:383-396; the CUDA tracer writes the linear value, selectable).
//   PFM : raw linear float RGB, bottom row first as the format requires.
// No zlib dependency: the PNG stream uses stored (uncompressed) deflate blocks.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/mcpt.h"
#include "context.h"

namespace {

uint32_t crc_table[256];
bool crc_ready = false;
uint32_t crc32(const uint8_t* p, size_t n, uint32_t c = 0) {
    if (!crc_ready) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t k = i;
            for (int j = 0; j < 8; ++j) k = (k & 1u) ? 0xEDB88320u ^ (k >> 1) : k >> 1;
            crc_table[i] = k;
        }
        crc_ready = true;
    }
    c = ~c;
    for (size_t i = 0; i < n; ++i) c = crc_table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
    return ~c;
}

void be32(std::vector<uint8_t>& v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back(static_cast<uint8_t>(x >> s)); }

void chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
    be32(out, static_cast<uint32_t>(data.size()));
    std::vector<uint8_t> body(type, type + 4);
    body.insert(body.end(), data.begin(), data.end());
    out.insert(out.end(), body.begin(), body.end());
    be32(out, crc32(body.data(), body.size()));
}

}  // namespace

extern "C" {

int mcpt_write_png(const char* path, const float* rgb, int width, int height, float gamma) {
    return mcpt::guarded([&]() -> int {
    if (!path || !rgb || width <= 0 || height <= 0 || !(gamma > 0)) return mcpt::fail(MCPT_ERR_INVALID, "mcpt_write_png: bad arguments");
    const size_t stride = 1 + 3 * static_cast<size_t>(width);
    std::vector<uint8_t> raw(stride * static_cast<size_t>(height));
    const float inv = 1.0f / gamma;
    for (int y = 0; y < height; ++y) {
        uint8_t* row = &raw[stride * static_cast<size_t>(y)];
        row[0] = 0;  // filter: none
        for (int x = 0; x < 3 * width; ++x) {
            float v = rgb[static_cast<size_t>(y) * 3 * width + x];
            v = v > 0 ? std::pow(v, inv) : 0.0f;
            v = v < 1 ? v : 1.0f;
            row[1 + x] = static_cast<uint8_t>(v * 255.0f + 0.5f);
        }
    }
    std::vector<uint8_t> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;  // adler32
    for (size_t off = 0; off < raw.size();) {
        const size_t n = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + n == raw.size() ? 1 : 0);
        z.push_back(static_cast<uint8_t>(n)); z.push_back(static_cast<uint8_t>(n >> 8));
        z.push_back(static_cast<uint8_t>(~n)); z.push_back(static_cast<uint8_t>((~n) >> 8));
        z.insert(z.end(), raw.begin() + static_cast<long>(off), raw.begin() + static_cast<long>(off + n));
        for (size_t i = off; i < off + n; ++i) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
        off += n;
    }
    be32(z, (b << 16) | a);
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    be32(ihdr, static_cast<uint32_t>(width)); be32(ihdr, static_cast<uint32_t>(height));
    ihdr.insert(ihdr.end(), {8, 2, 0, 0, 0});
    chunk(out, "IHDR", ihdr);
    chunk(out, "IDAT", z);
    chunk(out, "IEND", {});
    FILE* f = std::fopen(path, "wb");
    if (!f) return mcpt::fail(MCPT_ERR_IO, std::string("cannot write ") + path);
    const bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    std::fclose(f);
    return ok ? MCPT_OK : mcpt::fail(MCPT_ERR_IO, std::string("short write to ") + path);
    });
}

int mcpt_write_pfm(const char* path, const float* rgb, int width, int height) {
    return mcpt::guarded([&]() -> int {
    if (!path || !rgb || width <= 0 || height <= 0) return mcpt::fail(MCPT_ERR_INVALID, "mcpt_write_pfm: bad arguments");
    FILE* f = std::fopen(path, "wb");
    if (!f) return mcpt::fail(MCPT_ERR_IO, std::string("cannot write ") + path);
    std::fprintf(f, "PF\n%d %d\n-1.0\n", width, height);
    bool ok = true;
    for (int y = height - 1; y >= 0 && ok; --y)
        ok = std::fwrite(rgb + static_cast<size_t>(y) * 3 * width, sizeof(float), 3 * static_cast<size_t>(width), f) == 3 * static_cast<size_t>(width);
    std::fclose(f);
    return ok ? MCPT_OK : mcpt::fail(MCPT_ERR_IO, std::string("short write to ") + path);
    });
}

}  // extern "C"
