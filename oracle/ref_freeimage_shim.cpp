// ref_freeimage_shim.cpp — stand-in for the FreeImage functions the reference calls (include/picture.h:25-44 and the
// never-called helpers of include/debug_utils.h:24-48); the reference ships only Windows binaries of FreeImage.
// Decoding goes through stb_image, whose implementation is already instantiated by the reference's src/mesh.cu:6-9.
// FreeImage's conventions are kept: rows bottom-up, BGR(A) byte order, pitch rounded up to 4 bytes.
// Test/benchmark infrastructure only (built into oracle/_ref/).
#include <cstdlib>
#include <cstring>
#include "FreeImage/FreeImage.h"

extern "C" {
unsigned char* stbi_load(char const* filename, int* x, int* y, int* channels_in_file, int desired_channels);
void stbi_image_free(void* retval_from_stbi_load);
int stbi_write_png(char const* filename, int w, int h, int comp, const void* data, int stride_in_bytes);
}

namespace {
struct Bitmap {
    unsigned w, h, pitch, bpp;
    unsigned char* bits;
};
Bitmap* B(FIBITMAP* d) { return static_cast<Bitmap*>(d->data); }
}

extern "C" {
FREE_IMAGE_FORMAT DLL_CALLCONV FreeImage_GetFIFFromFormat(const char*) { return FIF_UNKNOWN; }   // stb sniffs the format itself

FIBITMAP* DLL_CALLCONV FreeImage_Allocate(int width, int height, int bpp, unsigned, unsigned, unsigned) {
    Bitmap* b = new Bitmap;
    b->w = (unsigned)width; b->h = (unsigned)height; b->bpp = (unsigned)bpp;
    b->pitch = ((b->w * (b->bpp / 8)) + 3u) & ~3u;
    b->bits = static_cast<unsigned char*>(calloc((size_t)b->pitch * b->h, 1));
    FIBITMAP* d = new FIBITMAP;
    d->data = b;
    return d;
}

FIBITMAP* DLL_CALLCONV FreeImage_Load(FREE_IMAGE_FORMAT, const char* filename, int) {
    int w = 0, h = 0, c = 0;
    unsigned char* px = stbi_load(filename, &w, &h, &c, 0);
    if (!px) return nullptr;
    if (c != 3 && c != 4) {   // FreeImage would hand back 8/16 bpp; the reference then throws (src/texture.cu:79-82)
        stbi_image_free(px);
        px = stbi_load(filename, &w, &h, &c, 3);
        c = 3;
        if (!px) return nullptr;
    }
    FIBITMAP* d = FreeImage_Allocate(w, h, c * 8, 0, 0, 0);
    Bitmap* b = B(d);
    for (int y = 0; y < h; ++y) {
        const unsigned char* src = px + (size_t)(h - 1 - y) * w * c;   // bottom-up
        unsigned char* dst = b->bits + (size_t)y * b->pitch;
        for (int x = 0; x < w; ++x) {
            dst[x * c + 0] = src[x * c + 2];   // B
            dst[x * c + 1] = src[x * c + 1];   // G
            dst[x * c + 2] = src[x * c + 0];   // R
            if (c == 4) dst[x * c + 3] = src[x * c + 3];
        }
    }
    stbi_image_free(px);
    return d;
}

unsigned DLL_CALLCONV FreeImage_GetWidth(FIBITMAP* d) { return B(d)->w; }
unsigned DLL_CALLCONV FreeImage_GetHeight(FIBITMAP* d) { return B(d)->h; }
unsigned DLL_CALLCONV FreeImage_GetPitch(FIBITMAP* d) { return B(d)->pitch; }
unsigned DLL_CALLCONV FreeImage_GetBPP(FIBITMAP* d) { return B(d)->bpp; }
BYTE* DLL_CALLCONV FreeImage_GetBits(FIBITMAP* d) { return B(d)->bits; }

void DLL_CALLCONV FreeImage_Unload(FIBITMAP* d) {
    if (!d) return;
    free(B(d)->bits);
    delete B(d);
    delete d;
}

BOOL DLL_CALLCONV FreeImage_Save(FREE_IMAGE_FORMAT, FIBITMAP* d, const char* filename, int) {
    Bitmap* b = B(d);
    const unsigned c = b->bpp / 8;
    unsigned char* rgb = static_cast<unsigned char*>(malloc((size_t)b->w * b->h * c));
    for (unsigned y = 0; y < b->h; ++y)
        for (unsigned x = 0; x < b->w; ++x) {
            const unsigned char* s = b->bits + (size_t)(b->h - 1 - y) * b->pitch + x * c;
            unsigned char* t = rgb + ((size_t)y * b->w + x) * c;
            t[0] = s[2]; t[1] = s[1]; t[2] = s[0];
            if (c == 4) t[3] = s[3];
        }
    int ok = stbi_write_png(filename, (int)b->w, (int)b->h, (int)c, rgb, (int)(b->w * c));
    free(rgb);
    return ok ? TRUE : FALSE;
}
}
