// Image output of the host front end: the writer is picked by file extension
// as in the reference's WriteImage (src/core/imageio.cpp:81-122).
//   .pfm  float32 RGB, bottom row first, scale -1 (little endian)   imageio.cpp:420-470
//   .exr  half RGB scanlines with dataWindow = cropped pixel bounds and
//         displayWindow = full resolution, as WriteImageEXR sets them
//         (imageio.cpp:164-189).  Written uncompressed (the reference's
//         RgbaOutputFile default is PIZ; readers accept either).
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "setup.h"

namespace pg {

static bool hasExt(const std::string &s, const char *ext) {
    size_t n = strlen(ext);
    if (s.size() < n) return false;
    for (size_t i = 0; i < n; ++i)
        if (tolower(s[s.size() - n + i]) != ext[i]) return false;
    return true;
}

static uint16_t floatToHalf(float f) {
    uint32_t x;
    memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    int32_t exp = (int32_t)((x >> 23) & 0xff) - 127 + 15;
    uint32_t mant = x & 0x7fffffu;
    if (((x >> 23) & 0xff) == 0xff) return (uint16_t)(sign | 0x7c00u | (mant ? 0x200u : 0));  // inf / nan
    if (exp >= 31) return (uint16_t)(sign | 0x7c00u);                                           // overflow -> inf
    if (exp <= 0) {
        if (exp < -10) return (uint16_t)sign;  // underflow -> 0
        mant |= 0x800000u;
        int shift = 14 - exp;
        uint32_t h = mant >> shift;
        uint32_t rem = mant & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) ++h;
        return (uint16_t)(sign | h);
    }
    uint32_t h = ((uint32_t)exp << 10) | (mant >> 13);
    uint32_t rem = mant & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;  // may carry into the exponent: correct
    return (uint16_t)(sign | h);
}

static bool writePFM(const std::string &name, const float *rgb, int w, int h) {
    FILE *f = fopen(name.c_str(), "wb");
    if (!f) {
        hostError("Unable to open output PFM file \"%s\"", name.c_str());
        return false;
    }
    fprintf(f, "PF\n%d %d\n-1.000000\n", w, h);
    for (int y = h - 1; y >= 0; --y) fwrite(rgb + (size_t)3 * w * y, sizeof(float), (size_t)3 * w, f);
    fclose(f);
    return true;
}

namespace {
struct ExrOut {
    std::vector<uint8_t> b;
    void str(const char *s) { b.insert(b.end(), s, s + strlen(s) + 1); }
    void i32(int32_t v) { uint8_t t[4]; memcpy(t, &v, 4); b.insert(b.end(), t, t + 4); }
    void f32(float v) { uint8_t t[4]; memcpy(t, &v, 4); b.insert(b.end(), t, t + 4); }
    void u64(uint64_t v) { uint8_t t[8]; memcpy(t, &v, 8); b.insert(b.end(), t, t + 8); }
    void attr(const char *name, const char *type, int32_t size) { str(name); str(type); i32(size); }
};
}  // namespace

static bool writeEXR(const std::string &name, const float *rgb, int w, int h, int totalX, int totalY, int xOff, int yOff) {
    ExrOut o;
    o.i32(20000630);  // magic
    o.i32(2);         // version 2, scanline, single part
    o.attr("channels", "chlist", 3 * 18 + 1);
    for (const char *c : {"B", "G", "R"}) {
        o.str(c);
        o.i32(1);  // HALF
        o.b.push_back(0); o.b.push_back(0); o.b.push_back(0); o.b.push_back(0);  // pLinear + reserved
        o.i32(1); o.i32(1);
    }
    o.b.push_back(0);
    o.attr("compression", "compression", 1); o.b.push_back(0);
    o.attr("dataWindow", "box2i", 16); o.i32(xOff); o.i32(yOff); o.i32(xOff + w - 1); o.i32(yOff + h - 1);
    o.attr("displayWindow", "box2i", 16); o.i32(0); o.i32(0); o.i32(totalX - 1); o.i32(totalY - 1);
    o.attr("lineOrder", "lineOrder", 1); o.b.push_back(0);
    o.attr("pixelAspectRatio", "float", 4); o.f32(1.f);
    o.attr("screenWindowCenter", "v2f", 8); o.f32(0.f); o.f32(0.f);
    o.attr("screenWindowWidth", "float", 4); o.f32(1.f);
    o.b.push_back(0);  // end of header
    size_t tableAt = o.b.size();
    size_t rowBytes = (size_t)w * 2 * 3, chunk = 8 + rowBytes;
    for (int y = 0; y < h; ++y) o.u64((uint64_t)(tableAt + (size_t)8 * h + chunk * y));
    std::vector<uint16_t> row((size_t)3 * w);
    for (int y = 0; y < h; ++y) {
        o.i32(yOff + y);
        o.i32((int32_t)rowBytes);
        for (int c = 0; c < 3; ++c)  // channel order B, G, R
            for (int x = 0; x < w; ++x) row[(size_t)c * w + x] = floatToHalf(rgb[3 * ((size_t)y * w + x) + (2 - c)]);
        const uint8_t *p = (const uint8_t *)row.data();
        o.b.insert(o.b.end(), p, p + rowBytes);
    }
    FILE *f = fopen(name.c_str(), "wb");
    if (!f) {
        hostError("Error writing \"%s\"", name.c_str());
        return false;
    }
    fwrite(o.b.data(), 1, o.b.size(), f);
    fclose(f);
    return true;
}

bool writeImage(const std::string &name, const float *rgb, const int cb[4], const int fullRes[2]) {
    int w = cb[2] - cb[0], h = cb[3] - cb[1];
    if (hasExt(name, ".exr")) return writeEXR(name, rgb, w, h, fullRes[0], fullRes[1], cb[0], cb[1]);
    if (hasExt(name, ".pfm")) return writePFM(name, rgb, w, h);
    hostError("Can't determine image file type from suffix of filename \"%s\" (this build writes .exr and .pfm)", name.c_str());
    return false;
}

bool readPFM(const std::string &name, std::vector<float> *rgb, int *w, int *h) {
    FILE *f = fopen(name.c_str(), "rb");
    if (!f) return false;
    char magic[3] = {0};
    float scale;
    if (fscanf(f, "%2s %d %d %f", magic, w, h, &scale) != 4 || strcmp(magic, "PF")) {
        fclose(f);
        return false;
    }
    fgetc(f);
    std::vector<float> rows((size_t)3 * *w * *h);
    size_t got = fread(rows.data(), sizeof(float), rows.size(), f);
    fclose(f);
    if (got != rows.size()) return false;
    rgb->resize(rows.size());
    for (int y = 0; y < *h; ++y)
        memcpy(rgb->data() + (size_t)3 * *w * y, rows.data() + (size_t)3 * *w * (*h - 1 - y), sizeof(float) * 3 * *w);
    return true;
}

}  // namespace pg
