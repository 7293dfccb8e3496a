// 8-bit RGB PNG output: the reference's imshow quantisation (MTPC/MTPC.cpp:22-30:
// (uchar)clamp(v*255, 0, 255), truncation, no gamma) followed by the uncompressed PNG stream
// MTPC/svpng.inc:77-108 emits byte by byte with fputc — the SAME bytes (one stored deflate block per
// scanline, tests/test_host.py compares with a file the reference wrote); here built in memory, written
// once, and the file is closed — the reference never fcloses, which truncated its veach-mis renders.
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "../../include/mcpt.h"

void mcpt_set_error(const char* msg);

namespace {
struct Crc {
    uint32_t table[256];
    Crc() {
        for (uint32_t n = 0; n < 256; n++) {
            uint32_t c = n;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[n] = c;
        }
    }
    uint32_t run(const uint8_t* p, size_t n, uint32_t c = 0xFFFFFFFFu) const {
        for (size_t i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
        return c;
    }
};
void be32(std::vector<uint8_t>& o, uint32_t v) { for (int s = 24; s >= 0; s -= 8) o.push_back(uint8_t(v >> s)); }
void chunk(std::vector<uint8_t>& o, const char tag[4], const std::vector<uint8_t>& body) {
    static const Crc crc;
    be32(o, uint32_t(body.size()));
    size_t at = o.size();
    o.insert(o.end(), tag, tag + 4);
    o.insert(o.end(), body.begin(), body.end());
    be32(o, ~crc.run(o.data() + at, o.size() - at));
}
}  // namespace

namespace mcpt {
void quantize_rgb(const float* rgb, size_t n, uint8_t* out) {
    for (size_t i = 0; i < n; i++) {
        double v = double(rgb[i]) * 255;
        v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);   // NaN falls through to 0 via the cast below
        out[i] = v == v ? uint8_t(v) : 0;
    }
}
}  // namespace mcpt

extern "C" MCPT_API int mcpt_write_png(const char* file, const float* rgb, int width, int height) {
    if (!file || !rgb || width <= 0 || height <= 0) { mcpt_set_error("mcpt_write_png: bad argument"); return MCPT_ERR_ARG; }
    // svpng stores a scanline's 3w+1 bytes in one deflate block whose length field has 16 bits (MTPC/svpng.inc:100)
    if (width > 21844) { mcpt_set_error("mcpt_write_png: width > 21844 does not fit svpng's one-block-per-scanline framing"); return MCPT_ERR_ARG; }
    const size_t row = size_t(width) * 3;
    // raw scanlines: filter byte 0 + RGB
    std::vector<uint8_t> raw((row + 1) * height);
    for (int y = 0; y < height; y++) {
        raw[(row + 1) * y] = 0;
        mcpt::quantize_rgb(rgb + row * y, row, &raw[(row + 1) * y + 1]);
    }
    // zlib container exactly as svpng frames it (MTPC/svpng.inc:96-105): ONE stored deflate block per scanline
    // (final-flag byte, LEN = 3w+1 and ~LEN little-endian, the filter byte 0, the row), IDAT length 2 + h*(5 + p) + 4
    std::vector<uint8_t> z;
    z.reserve(raw.size() + size_t(height) * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    const uint32_t p = uint32_t(row + 1);
    for (int y = 0; y < height; y++) {
        z.push_back(y == height - 1 ? 1 : 0);
        z.push_back(uint8_t(p)); z.push_back(uint8_t(p >> 8));
        z.push_back(uint8_t(~p)); z.push_back(uint8_t((~p) >> 8));
        const uint8_t* line = &raw[size_t(p) * y];
        z.insert(z.end(), line, line + p);
        for (uint32_t i = 0; i < p; i++) { a = (a + line[i]) % 65521; b = (b + a) % 65521; }
    }
    be32(z, (b << 16) | a);
    std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
    std::vector<uint8_t> ihdr;
    be32(ihdr, uint32_t(width)); be32(ihdr, uint32_t(height));
    ihdr.insert(ihdr.end(), {8, 2, 0, 0, 0});
    chunk(png, "IHDR", ihdr);
    chunk(png, "IDAT", z);
    chunk(png, "IEND", {});
    FILE* f = std::fopen(file, "wb");
    if (!f) { mcpt_set_error((std::string("cannot open ") + file).c_str()); return MCPT_ERR_IO; }
    bool ok = std::fwrite(png.data(), 1, png.size(), f) == png.size();
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { mcpt_set_error("short write"); return MCPT_ERR_IO; }
    return MCPT_OK;
}
