// PNG container + zlib for the slice files (see png_io.hpp).  Host code; nothing here is on the reward path.
#include "png_io.hpp"

#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace icr_png {
namespace {

const uint8_t SIG[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};

void put32(std::vector<uint8_t>& o, uint32_t v) {
    o.push_back(v >> 24); o.push_back(v >> 16); o.push_back(v >> 8); o.push_back(v);
}
uint32_t get32(const uint8_t* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }

void chunk(std::vector<uint8_t>& o, const char type[4], const uint8_t* data, size_t n) {
    put32(o, (uint32_t)n);
    const size_t at = o.size();
    o.insert(o.end(), type, type + 4);
    if (n) o.insert(o.end(), data, data + n);
    put32(o, (uint32_t)crc32(0L, o.data() + at, (uInt)(n + 4)));
}

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

}  // namespace

std::vector<uint8_t> encode_rgba(const uint8_t* rgba, int w, int h, int level) {
    const size_t stride = (size_t)w * 4;
    std::vector<uint8_t> raw((stride + 1) * h);
    for (int y = 0; y < h; ++y) {
        uint8_t* d = raw.data() + (stride + 1) * y;
        const uint8_t* s = rgba + stride * y;
        *d++ = 1;                                                    // Sub
        for (size_t i = 0; i < stride; ++i) d[i] = (uint8_t)(s[i] - (i >= 4 ? s[i - 4] : 0));
    }
    z_stream z;
    std::memset(&z, 0, sizeof z);
    deflateInit2(&z, level, Z_DEFLATED, 15, 8, Z_DEFAULT_STRATEGY);
    std::vector<uint8_t> comp(deflateBound(&z, (uLong)raw.size()));
    z.next_in = raw.data(); z.avail_in = (uInt)raw.size();
    z.next_out = comp.data(); z.avail_out = (uInt)comp.size();
    deflate(&z, Z_FINISH);
    comp.resize(z.total_out);
    deflateEnd(&z);

    std::vector<uint8_t> o(SIG, SIG + 8);
    uint8_t ihdr[13];
    const uint32_t uw = (uint32_t)w, uh = (uint32_t)h;
    ihdr[0] = uw >> 24; ihdr[1] = uw >> 16; ihdr[2] = uw >> 8; ihdr[3] = uw;
    ihdr[4] = uh >> 24; ihdr[5] = uh >> 16; ihdr[6] = uh >> 8; ihdr[7] = uh;
    ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;      // 8-bit RGBA, no interlace
    chunk(o, "IHDR", ihdr, 13);
    chunk(o, "IDAT", comp.data(), comp.size());
    chunk(o, "IEND", nullptr, 0);
    return o;
}

bool write_file(const std::string& path, const std::vector<uint8_t>& bytes) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    return std::fclose(f) == 0 && ok;
}

bool read_file(const std::string& path, std::vector<uint8_t>& bytes) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    bytes.resize(n > 0 ? (size_t)n : 0);
    const bool ok = n >= 0 && std::fread(bytes.data(), 1, bytes.size(), f) == bytes.size();
    std::fclose(f);
    return ok;
}

bool decode_rgba(const std::vector<uint8_t>& file, std::vector<uint8_t>& rgba, int& w, int& h, std::string& err) {
    if (file.size() < 8 + 25 || std::memcmp(file.data(), SIG, 8) != 0) { err = "not a PNG file"; return false; }
    size_t at = 8;
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat;
    w = h = 0;
    while (at + 12 <= file.size()) {
        const uint32_t n = get32(&file[at]);
        if (at + 12 + (size_t)n > file.size()) { err = "truncated chunk"; return false; }
        const uint8_t* type = &file[at + 4];
        const uint8_t* data = &file[at + 8];
        if (!std::memcmp(type, "IHDR", 4) && n >= 13) {
            w = (int)get32(data); h = (int)get32(data + 4);
            depth = data[8]; ctype = data[9]; interlace = data[12];
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), data, data + n);
        } else if (!std::memcmp(type, "IEND", 4)) {
            break;
        }
        at += 12 + (size_t)n;
    }
    int ch = ctype == 6 ? 4 : ctype == 2 ? 3 : ctype == 4 ? 2 : ctype == 0 ? 1 : 0;
    if (w <= 0 || h <= 0 || depth != 8 || ch == 0 || interlace != 0) { err = "unsupported PNG (need 8-bit grey/RGB/RGBA, no interlace)"; return false; }
    const size_t stride = (size_t)w * ch;
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf got = (uLongf)raw.size();
    if (uncompress(raw.data(), &got, idat.data(), (uLong)idat.size()) != Z_OK || got != raw.size()) { err = "bad IDAT stream"; return false; }
    std::vector<uint8_t> img(stride * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = raw.data() + (stride + 1) * y;
        const int f = *s++;
        uint8_t* d = img.data() + stride * y;
        const uint8_t* up = y ? d - stride : nullptr;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= (size_t)ch ? d[i - ch] : 0, b = up ? up[i] : 0, c = (up && i >= (size_t)ch) ? up[i - ch] : 0;
            int pred;
            switch (f) {
                case 0: pred = 0; break;
                case 1: pred = a; break;
                case 2: pred = b; break;
                case 3: pred = (a + b) >> 1; break;
                case 4: pred = paeth(a, b, c); break;
                default: err = "bad filter type"; return false;
            }
            d[i] = (uint8_t)(s[i] + pred);
        }
    }
    rgba.resize((size_t)w * h * 4);
    for (size_t p = 0; p < (size_t)w * h; ++p) {
        const uint8_t* s = &img[p * ch];
        uint8_t* d = &rgba[p * 4];
        if (ch == 4) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = s[3]; }
        else if (ch == 3) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = 255; }
        else if (ch == 2) { d[0] = d[1] = d[2] = s[0]; d[3] = s[1]; }
        else { d[0] = d[1] = d[2] = s[0]; d[3] = 255; }
    }
    return true;
}

}  // namespace icr_png
