// image_io.cpp — 8-bit image codecs for the host layer: PNG (zlib) and binary PPM in, PNG out.
//
// The reference loads its environment map and (in never-called debug helpers) saves PNGs through FreeImage
// (include/picture.h:19-45, include/debug_utils.h:19-49); FreeImage is a Windows binary in that tree.  These are
// this repository's own codecs: non-interlaced 8-bit grey/RGB/RGBA/palette PNG, P6 PPM.  JPEG is not decoded —
// convert the environment map to PNG first.
#include <zlib.h>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

#include "tpt/host.hpp"

namespace tpt {

static std::string slurp(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Picture: cannot open " + path);
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}
static uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

static void loadPPM(const std::string& data, Picture& pic) {
    std::istringstream in(data);
    std::string magic;
    int maxv = 0;
    in >> magic;
    auto skip = [&]() {
        while (true) {
            in >> std::ws;
            if (in.peek() == '#') { std::string l; std::getline(in, l); } else break;
        }
    };
    skip(); in >> pic.width; skip(); in >> pic.height; skip(); in >> maxv;
    in.get();
    if (magic != "P6" || maxv != 255 || pic.width <= 0 || pic.height <= 0) throw std::runtime_error("Picture: only 8-bit P6 PPM is supported");
    pic.channels = 3;
    pic.pixels.resize((size_t)pic.width * pic.height * 3);
    in.read(reinterpret_cast<char*>(pic.pixels.data()), (std::streamsize)pic.pixels.size());
    if ((size_t)in.gcount() != pic.pixels.size()) throw std::runtime_error("Picture: truncated PPM");
}

static void loadPNG(const std::string& data, Picture& pic) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(data.data());
    const size_t n = data.size();
    size_t off = 8;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, palette;
    while (off + 12 <= n) {
        const uint32_t len = be32(p + off);
        const char* type = reinterpret_cast<const char*>(p + off + 4);
        const uint8_t* body = p + off + 8;
        if (off + 12 + len > n) throw std::runtime_error("Picture: truncated PNG chunk");
        if (!std::memcmp(type, "IHDR", 4)) {
            pic.width = (int)be32(body);
            pic.height = (int)be32(body + 4);
            depth = body[8]; ctype = body[9]; interlace = body[12];
        } else if (!std::memcmp(type, "PLTE", 4)) {
            palette.assign(body, body + len);
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            break;
        }
        off += 12 + len;
    }
    if (depth != 8 || interlace != 0 || pic.width <= 0 || pic.height <= 0) throw std::runtime_error("Picture: only non-interlaced 8-bit PNG is supported");
    const int inCh = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!inCh) throw std::runtime_error("Picture: unsupported PNG colour type");
    const size_t stride = (size_t)pic.width * inCh;
    std::vector<uint8_t> raw((stride + 1) * pic.height);
    uLongf rawLen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rawLen, idat.data(), (uLong)idat.size()) != Z_OK || rawLen != raw.size())
        throw std::runtime_error("Picture: PNG inflate failed");
    std::vector<uint8_t> img(stride * pic.height);
    for (int y = 0; y < pic.height; ++y) {   // undo the scanline filters (PNG spec 9.2)
        const uint8_t ft = raw[(stride + 1) * y];
        const uint8_t* src = &raw[(stride + 1) * y + 1];
        uint8_t* dst = &img[stride * y];
        const uint8_t* up = y ? &img[stride * (y - 1)] : nullptr;
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= (size_t)inCh ? dst[x - inCh] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)inCh) ? up[x - inCh] : 0;
            int pred = 0;
            switch (ft) {
                case 0: pred = 0; break;
                case 1: pred = a; break;
                case 2: pred = b; break;
                case 3: pred = (a + b) >> 1; break;
                case 4: {
                    const int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c);
                    pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                    break;
                }
                default: throw std::runtime_error("Picture: bad PNG filter");
            }
            dst[x] = (uint8_t)(src[x] + pred);
        }
    }
    pic.channels = (ctype == 6 || ctype == 4) ? 4 : 3;
    pic.pixels.resize((size_t)pic.width * pic.height * pic.channels);
    for (size_t i = 0; i < (size_t)pic.width * pic.height; ++i) {
        uint8_t* o = &pic.pixels[i * pic.channels];
        const uint8_t* s = &img[i * inCh];
        switch (ctype) {
            case 0: o[0] = o[1] = o[2] = s[0]; break;
            case 2: o[0] = s[0]; o[1] = s[1]; o[2] = s[2]; break;
            case 3:
                if ((size_t)s[0] * 3 + 2 >= palette.size()) throw std::runtime_error("Picture: PNG palette index out of range");
                o[0] = palette[s[0] * 3]; o[1] = palette[s[0] * 3 + 1]; o[2] = palette[s[0] * 3 + 2];
                break;
            case 4: o[0] = o[1] = o[2] = s[0]; o[3] = s[1]; break;
            case 6: o[0] = s[0]; o[1] = s[1]; o[2] = s[2]; o[3] = s[3]; break;
        }
    }
}

Picture::Picture(const std::string& filename) {
    const std::string data = slurp(filename);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (data.size() > 8 && !std::memcmp(data.data(), sig, 8)) loadPNG(data, *this);
    else if (data.size() > 2 && data[0] == 'P' && data[1] == '6') loadPPM(data, *this);
    else throw std::runtime_error("Picture: " + filename + " is neither PNG nor binary PPM (JPEG is not decoded here: convert it first)");
    if (channels != 3 && channels != 4) throw std::runtime_error("Unsupported image format");   // src/texture.cu:79-82
}

static void chunk(std::vector<uint8_t>& out, const char* type, const uint8_t* body, size_t len) {
    const uint8_t l[4] = {(uint8_t)(len >> 24), (uint8_t)(len >> 16), (uint8_t)(len >> 8), (uint8_t)len};
    out.insert(out.end(), l, l + 4);
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    if (len) out.insert(out.end(), body, body + len);
    const uint32_t c = (uint32_t)crc32(0L, out.data() + start, (uInt)(len + 4));
    const uint8_t cb[4] = {(uint8_t)(c >> 24), (uint8_t)(c >> 16), (uint8_t)(c >> 8), (uint8_t)c};
    out.insert(out.end(), cb, cb + 4);
}

void writePNG(const std::string& filename, const uint8_t* px, int width, int height, int channels) {
    if (channels != 3 && channels != 4) throw std::runtime_error("writePNG: 3 or 4 channels");
    const size_t stride = (size_t)width * channels;
    std::vector<uint8_t> raw((stride + 1) * height);
    for (int y = 0; y < height; ++y) {
        raw[(stride + 1) * y] = 0;   // filter: none
        std::memcpy(&raw[(stride + 1) * y + 1], px + stride * y, stride);
    }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) throw std::runtime_error("writePNG: deflate failed");
    std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    uint8_t ihdr[13] = {(uint8_t)(width >> 24), (uint8_t)(width >> 16), (uint8_t)(width >> 8), (uint8_t)width,
                        (uint8_t)(height >> 24), (uint8_t)(height >> 16), (uint8_t)(height >> 8), (uint8_t)height,
                        8, (uint8_t)(channels == 4 ? 6 : 2), 0, 0, 0};
    chunk(out, "IHDR", ihdr, 13);
    chunk(out, "IDAT", comp.data(), clen);
    chunk(out, "IEND", nullptr, 0);
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f) throw std::runtime_error("writePNG: cannot create " + filename);
    std::fwrite(out.data(), 1, out.size(), f);
    std::fclose(f);
}

}   // namespace tpt
