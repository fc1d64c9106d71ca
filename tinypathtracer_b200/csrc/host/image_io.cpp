// image_io.cpp — 8-bit image codecs for the host layer: PNG (zlib) and binary PPM in, PNG out.
//
// The reference loads its environment map and (in never-called debug helpers) saves PNGs through FreeImage
// (include/picture.h:19-45, include/debug_utils.h:19-49); FreeImage is a Windows binary in that tree.  These are
// this repository's own codecs: non-interlaced 8-bit grey/RGB/RGBA/palette PNG, P6 PPM, baseline JPEG.
#include <zlib.h>

#include <algorithm>
#include <cmath>
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

// ---- JPEG: baseline / extended sequential DCT, 8-bit, Huffman (ITU T.81) ------------------------------------------
// Grey or YCbCr (JFIF) with any sampling factors, restart intervals; chroma is upsampled by replication.  Progressive,
// arithmetic-coded, lossless and 12-bit files are refused with a message.  The reference reads its environment map
// (a .jpg in its README and main.cu) through FreeImage; this is the decoder that stands in for it here.
namespace {
struct JpegHuff {
    uint8_t bits[17] = {0};
    uint8_t vals[256] = {0};
    int mincode[17] = {0}, maxcode[18] = {0}, valptr[17] = {0};
    bool defined = false;           // set by a DHT segment; a scan that names an undefined table is refused
    void build() {
        defined = true;
        int code = 0, k = 0;
        for (int l = 1; l <= 16; ++l) {
            valptr[l] = k;
            mincode[l] = code;
            code += bits[l];
            k += bits[l];
            maxcode[l] = bits[l] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7fffffff;
    }
};
struct JpegComp {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0, pred = 0;
    int bw = 0, bh = 0;             // size in blocks (padded to whole MCUs)
    std::vector<uint8_t> plane;     // bw*8 x bh*8 samples
};
struct JpegBits {
    const uint8_t* p;
    const uint8_t* end;
    uint32_t acc = 0;
    int n = 0;
    bool marker = false;            // ran into a marker (RSTn / EOI): feed zeros until the caller deals with it
    int bit() {
        if (n == 0) {
            uint8_t b = 0;
            if (!marker && p < end) {
                b = *p++;
                if (b == 0xFF) {
                    if (p < end && *p == 0x00) ++p;          // stuffed zero
                    else { marker = true; --p; b = 0; }      // leave the marker for restart()
                }
            }
            acc = b;
            n = 8;
        }
        --n;
        return (int)((acc >> n) & 1u);
    }
    int receive(int s) {
        int v = 0;
        for (int i = 0; i < s; ++i) v = (v << 1) | bit();
        return v;
    }
    void restart() {   // byte-align, skip the RSTn marker
        n = 0;
        marker = false;
        while (p + 1 < end && !(p[0] == 0xFF && p[1] >= 0xD0 && p[1] <= 0xD7)) ++p;
        if (p + 1 < end) p += 2;
    }
};
int jpegDecode(JpegBits& br, const JpegHuff& h) {
    int code = 0;
    for (int l = 1; l <= 16; ++l) {
        code = (code << 1) | br.bit();
        if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) return h.vals[h.valptr[l] + code - h.mincode[l]];
    }
    throw std::runtime_error("Picture: corrupt JPEG Huffman code");
}
inline int jpegExtend(int v, int s) { return (s && v < (1 << (s - 1))) ? v - (1 << s) + 1 : v; }
const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
void jpegIdct(const float* in, uint8_t* out, size_t stride) {   // separable float IDCT, +128, clamp
    static float c[8][8];
    static bool init = false;
    if (!init) {
        for (int x = 0; x < 8; ++x)
            for (int u = 0; u < 8; ++u) c[x][u] = (u == 0 ? 0.35355339059f : 0.5f) * std::cos((2 * x + 1) * u * 3.14159265358979323846 / 16.0);
        init = true;
    }
    float tmp[64];
    for (int v = 0; v < 8; ++v)
        for (int x = 0; x < 8; ++x) {
            float s = 0;
            for (int u = 0; u < 8; ++u) s += c[x][u] * in[v * 8 + u];
            tmp[v * 8 + x] = s;
        }
    for (int x = 0; x < 8; ++x)
        for (int y = 0; y < 8; ++y) {
            float s = 0;
            for (int v = 0; v < 8; ++v) s += c[y][v] * tmp[v * 8 + x];
            const int q = (int)std::lround(s + 128.0f);
            out[(size_t)y * stride + x] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
        }
}
}   // namespace

static void loadJPEG(const std::string& data, Picture& pic) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(data.data());
    const size_t n = data.size();
    uint16_t qt[4][64] = {};
    bool qtDefined[4] = {false, false, false, false};
    JpegHuff dc[4], ac[4];
    std::vector<JpegComp> comps;
    int restartInterval = 0, adobeTransform = -1;
    bool haveFrame = false;
    size_t off = 2;
    auto be16 = [&](size_t o) { return (int)((p[o] << 8) | p[o + 1]); };
    while (off + 4 <= n) {
        if (p[off] != 0xFF) { ++off; continue; }
        const int m = p[off + 1];
        if (m == 0xFF) { ++off; continue; }              // fill byte
        if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { off += 2; continue; }
        if (m == 0xD9) break;
        const size_t len = (size_t)be16(off + 2);
        if (len < 2 || off + 2 + len > n) throw std::runtime_error("Picture: truncated JPEG segment");
        const uint8_t* seg = p + off + 4;
        const size_t segLen = len - 2;
        if (m == 0xDB) {   // DQT
            size_t o = 0;
            while (o < segLen) {
                const int pq = seg[o] >> 4, tq = seg[o] & 15;
                ++o;
                if (tq > 3 || o + (pq ? 128 : 64) > segLen) throw std::runtime_error("Picture: bad JPEG quantisation table");
                for (int i = 0; i < 64; ++i) {
                    qt[tq][kZigzag[i]] = pq ? (uint16_t)((seg[o] << 8) | seg[o + 1]) : seg[o];
                    o += pq ? 2 : 1;
                }
                qtDefined[tq] = true;
            }
        } else if (m == 0xC4) {   // DHT
            size_t o = 0;
            while (o + 17 <= segLen) {
                const int tc = seg[o] >> 4, th = seg[o] & 15;
                if (th > 3 || tc > 1) throw std::runtime_error("Picture: bad JPEG Huffman table");
                JpegHuff& h = tc ? ac[th] : dc[th];
                int total = 0;
                for (int l = 1; l <= 16; ++l) { h.bits[l] = seg[o + l]; total += h.bits[l]; }
                o += 17;
                if (total > 256 || o + total > segLen) throw std::runtime_error("Picture: bad JPEG Huffman table");
                std::memcpy(h.vals, seg + o, (size_t)total);
                o += (size_t)total;
                h.build();
            }
        } else if (m == 0xC0 || m == 0xC1) {   // SOF0 / SOF1
            if (segLen < 6 || seg[0] != 8) throw std::runtime_error("Picture: only 8-bit JPEG is supported");
            pic.height = be16(off + 5);
            pic.width = be16(off + 7);
            const int nc = seg[5];
            if ((nc != 1 && nc != 3) || segLen < (size_t)(6 + 3 * nc) || pic.width <= 0 || pic.height <= 0)
                throw std::runtime_error("Picture: unsupported JPEG component count or size");
            comps.resize((size_t)nc);
            for (int i = 0; i < nc; ++i) {
                comps[i].id = seg[6 + 3 * i];
                comps[i].h = seg[7 + 3 * i] >> 4;
                comps[i].v = seg[7 + 3 * i] & 15;
                comps[i].tq = seg[8 + 3 * i] & 3;
                if (comps[i].h < 1 || comps[i].h > 4 || comps[i].v < 1 || comps[i].v > 4) throw std::runtime_error("Picture: bad JPEG sampling factors");
            }
            haveFrame = true;
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
            throw std::runtime_error(m == 0xC2 ? "Picture: progressive JPEG is not supported (re-save as baseline or PNG)"
                                               : "Picture: this JPEG coding process is not supported (baseline Huffman only)");
        } else if (m == 0xDD) {
            restartInterval = be16(off + 4);
        } else if (m == 0xEE && segLen >= 12 && !std::memcmp(seg, "Adobe", 5)) {
            adobeTransform = seg[11];
        } else if (m == 0xDA) {   // SOS: one interleaved scan with every component
            if (!haveFrame) throw std::runtime_error("Picture: JPEG scan before frame header");
            if (segLen < 1) throw std::runtime_error("Picture: truncated JPEG scan header");
            const int ns = seg[0];
            if (ns != (int)comps.size()) throw std::runtime_error("Picture: non-interleaved JPEG scans are not supported");
            if (segLen < (size_t)(1 + 2 * ns + 3)) throw std::runtime_error("Picture: truncated JPEG scan header");
            for (int i = 0; i < ns; ++i) {
                const int cid = seg[1 + 2 * i];
                const int td = seg[2 + 2 * i] >> 4, ta = seg[2 + 2 * i] & 15;
                if (td > 3 || ta > 3) throw std::runtime_error("Picture: JPEG scan names a Huffman table above 3");
                for (JpegComp& c : comps)
                    if (c.id == cid) { c.td = td; c.ta = ta; }
            }
            for (const JpegComp& c : comps) {
                if (!dc[c.td].defined || !ac[c.ta].defined) throw std::runtime_error("Picture: JPEG scan uses a Huffman table no DHT segment defined");
                if (!qtDefined[c.tq]) throw std::runtime_error("Picture: JPEG component uses a quantisation table no DQT segment defined");
            }
            int hmax = 1, vmax = 1;
            for (const JpegComp& c : comps) { hmax = std::max(hmax, c.h); vmax = std::max(vmax, c.v); }
            const int mcuW = 8 * hmax, mcuH = 8 * vmax;
            const int mcusX = (pic.width + mcuW - 1) / mcuW, mcusY = (pic.height + mcuH - 1) / mcuH;
            for (JpegComp& c : comps) {
                c.bw = mcusX * c.h;
                c.bh = mcusY * c.v;
                c.plane.assign((size_t)c.bw * 8 * c.bh * 8, 0);
                c.pred = 0;
            }
            JpegBits br{p + off + 2 + len, p + n};
            float blk[64];
            int untilRestart = restartInterval;
            for (int my = 0; my < mcusY; ++my)
                for (int mx = 0; mx < mcusX; ++mx) {
                    if (restartInterval && untilRestart == 0) {
                        br.restart();
                        for (JpegComp& c : comps) c.pred = 0;
                        untilRestart = restartInterval;
                    }
                    for (JpegComp& c : comps)
                        for (int by = 0; by < c.v; ++by)
                            for (int bx = 0; bx < c.h; ++bx) {
                                std::memset(blk, 0, sizeof blk);
                                const int t = jpegDecode(br, dc[c.td]);
                                if (t > 16) throw std::runtime_error("Picture: corrupt JPEG DC size category");
                                c.pred += jpegExtend(br.receive(t), t);
                                blk[0] = (float)(c.pred * (int)qt[c.tq][0]);
                                for (int k = 1; k < 64;) {
                                    const int rs = jpegDecode(br, ac[c.ta]), r = rs >> 4, sz = rs & 15;
                                    if (sz == 0) {
                                        if (r != 15) break;   // EOB
                                        k += 16;
                                        continue;
                                    }
                                    k += r;
                                    if (k > 63) throw std::runtime_error("Picture: corrupt JPEG block");
                                    const int zz = kZigzag[k];
                                    blk[zz] = (float)(jpegExtend(br.receive(sz), sz) * (int)qt[c.tq][zz]);
                                    ++k;
                                }
                                const size_t stride = (size_t)c.bw * 8;
                                jpegIdct(blk, &c.plane[((size_t)(my * c.v + by) * 8) * stride + (size_t)(mx * c.h + bx) * 8], stride);
                            }
                    --untilRestart;
                }
            // Subsampled components -> full resolution.  2:1 horizontally (and 2:1 vertically) uses libjpeg's "fancy"
            // triangle filter (jdsample.c h2v1 / h2v2, its default), so the result matches what FreeImage/libjpeg
            // readers give; other ratios replicate.
            std::vector<std::vector<uint8_t>> full(comps.size());
            for (size_t ci = 0; ci < comps.size(); ++ci) {
                const JpegComp& c = comps[ci];
                const int fx = hmax / c.h, fy = vmax / c.v;
                const size_t ps = (size_t)c.bw * 8;
                const int dw = (pic.width * c.h + hmax - 1) / hmax, dh = (pic.height * c.v + vmax - 1) / vmax;
                std::vector<uint8_t>& out = full[ci];
                out.assign((size_t)pic.width * pic.height, 0);
                const bool fancy = hmax % c.h == 0 && vmax % c.v == 0 && fx == 2 && (fy == 1 || fy == 2) && dw >= 2;
                if (!fancy) {
                    for (int y = 0; y < pic.height; ++y)
                        for (int x = 0; x < pic.width; ++x) out[(size_t)y * pic.width + x] = c.plane[(size_t)(y * c.v / vmax) * ps + (size_t)(x * c.h / hmax)];
                    continue;
                }
                std::vector<int> col((size_t)dw);
                for (int y = 0; y < pic.height; ++y) {
                    int round0, round1, shift;
                    if (fy == 1) {   // h2v1: column sums are the samples themselves
                        const uint8_t* r0 = &c.plane[(size_t)y * ps];
                        for (int i = 0; i < dw; ++i) col[(size_t)i] = r0[i];
                        round0 = 1; round1 = 2; shift = 2;
                    } else {         // h2v2: 3/4 of the nearer row + 1/4 of the next nearer one
                        const int sy = y >> 1, ny = (y & 1) ? std::min(sy + 1, dh - 1) : std::max(sy - 1, 0);
                        const uint8_t *r0 = &c.plane[(size_t)sy * ps], *r1 = &c.plane[(size_t)ny * ps];
                        for (int i = 0; i < dw; ++i) col[(size_t)i] = 3 * r0[i] + r1[i];
                        round0 = 8; round1 = 7; shift = 4;
                    }
                    const int unit = fy == 1 ? 1 : 4;   // weight of a column sum that stands alone at the image border
                    uint8_t* o = &out[(size_t)y * pic.width];
                    for (int x = 0; x < pic.width; ++x) {
                        const int i = x >> 1;
                        int v;
                        if ((x & 1) == 0) v = i == 0 ? (fy == 1 ? col[0] : (col[0] * unit + round0) >> shift) : (3 * col[(size_t)i] + col[(size_t)i - 1] + round0) >> shift;
                        else v = i == dw - 1 ? (fy == 1 ? col[(size_t)i] : (col[(size_t)i] * unit + round1) >> shift) : (3 * col[(size_t)i] + col[(size_t)i + 1] + round1) >> shift;
                        o[x] = (uint8_t)v;
                    }
                }
            }
            pic.channels = 3;
            pic.pixels.resize((size_t)pic.width * pic.height * 3);
            const bool ycc = comps.size() == 3 && adobeTransform != 0;
            for (int y = 0; y < pic.height; ++y)
                for (int x = 0; x < pic.width; ++x) {
                    int v[3] = {0, 128, 128};
                    for (size_t ci = 0; ci < comps.size(); ++ci) v[ci] = full[ci][(size_t)y * pic.width + x];
                    uint8_t* o = &pic.pixels[((size_t)y * pic.width + x) * 3];
                    if (comps.size() == 1) {
                        o[0] = o[1] = o[2] = (uint8_t)v[0];
                    } else if (ycc) {
                        const float Y = (float)v[0], cb = (float)v[1] - 128.0f, cr = (float)v[2] - 128.0f;
                        const float rgb[3] = {Y + 1.402f * cr, Y - 0.344136f * cb - 0.714136f * cr, Y + 1.772f * cb};
                        for (int k = 0; k < 3; ++k) {
                            const int q = (int)std::lround(rgb[k]);
                            o[k] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
                        }
                    } else {
                        o[0] = (uint8_t)v[0]; o[1] = (uint8_t)v[1]; o[2] = (uint8_t)v[2];
                    }
                }
            return;
        }
        off += 2 + len;
    }
    throw std::runtime_error("Picture: JPEG without image data");
}

Picture::Picture(const std::string& filename) {
    const std::string data = slurp(filename);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (data.size() > 8 && !std::memcmp(data.data(), sig, 8)) loadPNG(data, *this);
    else if (data.size() > 2 && data[0] == 'P' && data[1] == '6') loadPPM(data, *this);
    else if (data.size() > 3 && (uint8_t)data[0] == 0xFF && (uint8_t)data[1] == 0xD8) loadJPEG(data, *this);
    else throw std::runtime_error("Picture: " + filename + " is not a PNG, JPEG or binary PPM file");
    if (channels != 3 && channels != 4) throw std::runtime_error("Unsupported image format");   // src/texture.cu:79-82
}

// ---- float images: Radiance RGBE (.hdr) and PFM ---------------------------------------------------------
// The reference reads environment maps through FreeImage into an 8-bit Texture only (include/picture.h:25-44,
// src/texture.cu:79-82).  RGBE -> float follows the rgbe.c convention FreeImage's HDR plugin uses: mantissa * 2^(e - 136),
// e = 0 -> black (no +0.5 bias).
static bool nextLine(const std::string& d, size_t& pos, std::string& line) {
    if (pos >= d.size()) return false;
    const size_t e = d.find('\n', pos);
    line = d.substr(pos, e == std::string::npos ? std::string::npos : e - pos);
    pos = e == std::string::npos ? d.size() : e + 1;
    if (!line.empty() && line.back() == '\r') line.pop_back();
    return true;
}

static void loadHDR(const std::string& d, PictureF& pic) {
    size_t pos = 0;
    std::string line;
    bool format = false;
    nextLine(d, pos, line);   // "#?RADIANCE" / "#?RGBE"
    while (nextLine(d, pos, line) && !line.empty())
        if (line.rfind("FORMAT=", 0) == 0) {
            if (line != "FORMAT=32-bit_rle_rgbe") throw std::runtime_error("PictureF: unsupported Radiance format " + line);
            format = true;
        }
    if (!format) throw std::runtime_error("PictureF: Radiance header without FORMAT=32-bit_rle_rgbe");
    if (!nextLine(d, pos, line)) throw std::runtime_error("PictureF: Radiance file without a resolution line");
    int h = 0, w = 0;
    if (std::sscanf(line.c_str(), "-Y %d +X %d", &h, &w) != 2 || w <= 0 || h <= 0 || w > 65535 || h > 65535)
        throw std::runtime_error("PictureF: only the standard '-Y h +X w' orientation is supported");
    pic.width = w;
    pic.height = h;
    pic.pixels.assign((size_t)w * h * 3, 0.0f);
    std::vector<uint8_t> row((size_t)w * 4);
    const uint8_t* p = reinterpret_cast<const uint8_t*>(d.data()) + pos;
    const uint8_t* end = reinterpret_cast<const uint8_t*>(d.data()) + d.size();
    auto need = [&](size_t n) { if ((size_t)(end - p) < n) throw std::runtime_error("PictureF: truncated Radiance file"); };
    for (int y = 0; y < h; ++y) {
        need(4);
        if (w >= 8 && w < 32768 && p[0] == 2 && p[1] == 2 && !(p[2] & 0x80)) {   // new RLE: each of the 4 channels run-length coded
            if (((int)p[2] << 8 | p[3]) != w) throw std::runtime_error("PictureF: Radiance scanline width mismatch");
            p += 4;
            for (int c = 0; c < 4; ++c)
                for (int x = 0; x < w;) {
                    need(1);
                    int n = *p++;
                    if (n > 128) {   // run
                        n -= 128;
                        need(1);
                        if (x + n > w) throw std::runtime_error("PictureF: Radiance run overflows its scanline");
                        const uint8_t v = *p++;
                        while (n--) row[4 * (size_t)x++ + c] = v;
                    } else {         // literals
                        if (n == 0 || x + n > w) throw std::runtime_error("PictureF: bad Radiance literal count");
                        need((size_t)n);
                        while (n--) row[4 * (size_t)x++ + c] = *p++;
                    }
                }
        } else {   // flat RGBE pixels (old-style repeat markers are not produced by any current writer)
            need((size_t)w * 4);
            std::memcpy(row.data(), p, (size_t)w * 4);
            p += (size_t)w * 4;
        }
        float* out = &pic.pixels[(size_t)y * w * 3];
        for (int x = 0; x < w; ++x) {
            const uint8_t* q = &row[4 * (size_t)x];
            const float f = q[3] ? std::ldexp(1.0f, (int)q[3] - 136) : 0.0f;
            out[3 * x + 0] = q[0] * f;
            out[3 * x + 1] = q[1] * f;
            out[3 * x + 2] = q[2] * f;
        }
    }
}

// PFM: "PF" (RGB) or "Pf" (grey), "w h", scale (negative = little endian), then rows BOTTOM to top.
static void loadPFM(const std::string& d, PictureF& pic) {
    int w = 0, h = 0, used = 0;
    float scale = 0.0f;
    char kind = 0;
    if (std::sscanf(d.c_str(), "P%c %d %d %f%n", &kind, &w, &h, &scale, &used) != 4 || (kind != 'F' && kind != 'f') || w <= 0 || h <= 0 ||
        scale == 0.0f)
        throw std::runtime_error("PictureF: bad PFM header");
    const size_t pos = (size_t)used + 1;   // exactly one whitespace byte after the scale
    const int ch = kind == 'F' ? 3 : 1;
    if (d.size() < pos + (size_t)w * h * ch * 4) throw std::runtime_error("PictureF: truncated PFM file");
    pic.width = w;
    pic.height = h;
    pic.pixels.resize((size_t)w * h * 3);
    const uint8_t* p = reinterpret_cast<const uint8_t*>(d.data()) + pos;
    const bool little = scale < 0.0f;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
            for (int c = 0; c < 3; ++c) {
                const uint8_t* q = p + 4 * (((size_t)(h - 1 - y) * w + x) * ch + (ch == 3 ? c : 0));
                const uint32_t bits = little ? ((uint32_t)q[0] | (uint32_t)q[1] << 8 | (uint32_t)q[2] << 16 | (uint32_t)q[3] << 24) : be32(q);
                float v;
                std::memcpy(&v, &bits, 4);
                pic.pixels[((size_t)y * w + x) * 3 + c] = v;
            }
}

bool PictureF::isFloatFile(const std::string& filename) {
    FILE* f = std::fopen(filename.c_str(), "rb");
    if (!f) return false;
    char m[2] = {0, 0};
    const size_t n = std::fread(m, 1, 2, f);
    std::fclose(f);
    return n == 2 && ((m[0] == '#' && m[1] == '?') || (m[0] == 'P' && (m[1] == 'F' || m[1] == 'f')));
}

PictureF::PictureF(const std::string& filename) {
    const std::string data = slurp(filename);
    if (data.size() > 2 && data[0] == '#' && data[1] == '?') loadHDR(data, *this);
    else if (data.size() > 2 && data[0] == 'P' && (data[1] == 'F' || data[1] == 'f')) loadPFM(data, *this);
    else throw std::runtime_error("PictureF: " + filename + " is not a Radiance .hdr or PFM file");
}

void writePFM(const std::string& filename, const float* rgbTopDown, int width, int height) {
    FILE* f = std::fopen(filename.c_str(), "wb");
    if (!f) throw std::runtime_error("writePFM: cannot create " + filename);
    std::fprintf(f, "PF\n%d %d\n-1.0\n", width, height);   // little endian
    for (int y = height - 1; y >= 0; --y) std::fwrite(rgbTopDown + (size_t)y * width * 3, sizeof(float), (size_t)width * 3, f);
    std::fclose(f);
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
