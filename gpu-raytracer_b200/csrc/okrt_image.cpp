/*
 * okrt_image.cpp — row f3 of SURVEY.md §8: what Application::saveScreenshot (source/Application/Application.cpp:661-674)
 * does after Okay::getCPUTextureData (source/DirectX/DX11.cpp:196-230) has brought the RGBA8 target to the host:
 * stbi_write_png("render.png", W, H, 4, data, W * 4).  stb_image_write is a third-party single header (deps/, not part of
 * this path's algorithm); what matters downstream is the decoded image, so this is a plain PNG encoder of its own:
 * 8-bit RGBA, filter 0 on every row, zlib stream of deflate blocks with a run-length matcher (distance = one pixel and
 * one row) over the fixed Huffman code.  Plus the raw progressive-state file (RGBA32F accumulation + frame count) used to
 * checkpoint / resume long accumulations and to merge sample shards offline.
 */
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "okrt.h"
#include "okrt_internal.h"

namespace {

uint32_t crcTable[256];
bool crcReady = false;
uint32_t crc32(uint32_t crc, const uint8_t* p, size_t n)
{
    if (!crcReady) {
        for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1; crcTable[i] = c; }
        crcReady = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = crcTable[(crc ^ p[i]) & 0xFFu] ^ (crc >> 8);
    return ~crc;
}

struct BitWriter {
    std::vector<uint8_t>& out; uint32_t acc = 0; int bits = 0;
    explicit BitWriter(std::vector<uint8_t>& o) : out(o) {}
    void put(uint32_t value, int count) { acc |= value << bits; bits += count; while (bits >= 8) { out.push_back((uint8_t)acc); acc >>= 8; bits -= 8; } }
    void putHuffman(uint32_t code, int count) { uint32_t r = 0; for (int i = 0; i < count; i++) r |= ((code >> (count - 1 - i)) & 1u) << i; put(r, count); }   /* codes go MSB first */
    void flush() { if (bits) { out.push_back((uint8_t)acc); acc = 0; bits = 0; } }
};

/* fixed Huffman code of RFC 1951 section 3.2.6 */
void putLiteral(BitWriter& w, uint32_t v)
{
    if (v < 144) w.putHuffman(0x30 + v, 8);
    else if (v < 256) w.putHuffman(0x190 + (v - 144), 9);
    else if (v < 280) w.putHuffman(v - 256, 7);
    else w.putHuffman(0xC0 + (v - 280), 8);
}
void putLength(BitWriter& w, uint32_t len)          /* 3..258 */
{
    static const uint16_t base[29] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
    static const uint8_t extra[29] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
    int k = 28; while (base[k] > len) k--;
    putLiteral(w, 257 + (uint32_t)k);
    if (extra[k]) w.put(len - base[k], extra[k]);
}
void putDistance(BitWriter& w, uint32_t dist)       /* 1..32768 */
{
    static const uint16_t base[30] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
    static const uint8_t extra[30] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
    int k = 29; while (base[k] > dist) k--;
    w.putHuffman((uint32_t)k, 5);
    if (extra[k]) w.put(dist - base[k], extra[k]);
}

void chunk(std::vector<uint8_t>& file, const char* type, const std::vector<uint8_t>& data)
{
    const uint32_t n = (uint32_t)data.size();
    const uint8_t len[4] = {(uint8_t)(n >> 24), (uint8_t)(n >> 16), (uint8_t)(n >> 8), (uint8_t)n};
    file.insert(file.end(), len, len + 4);
    const size_t at = file.size();
    file.insert(file.end(), type, type + 4);
    file.insert(file.end(), data.begin(), data.end());
    const uint32_t c = crc32(0, file.data() + at, file.size() - at);
    const uint8_t crc[4] = {(uint8_t)(c >> 24), (uint8_t)(c >> 16), (uint8_t)(c >> 8), (uint8_t)c};
    file.insert(file.end(), crc, crc + 4);
}

} // namespace

/* Application.cpp:671 — stbi_write_png(path, W, H, 4, data, W * 4) */
extern "C" OKRT_API int okrt_write_png(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height)
{
    using okrt::setError;
    if (!path || !rgba8 || !width || !height) { setError("okrt_write_png: null or empty argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    const size_t rowBytes = (size_t)width * 4, stride = rowBytes + 1;
    std::vector<uint8_t> raw(stride * height);
    for (uint32_t y = 0; y < height; y++) {
        raw[y * stride] = 0;                                      /* filter type None */
        std::memcpy(&raw[y * stride + 1], rgba8 + (size_t)y * rowBytes, rowBytes);
    }
    uint32_t a = 1, b = 0;                                        /* Adler-32 of the unfiltered stream */
    for (size_t i = 0; i < raw.size();) {
        const size_t n = std::min<size_t>(raw.size() - i, 5552);
        for (size_t k = 0; k < n; k++) { a += raw[i + k]; b += a; }
        a %= 65521u; b %= 65521u; i += n;
    }
    std::vector<uint8_t> z;
    z.reserve(raw.size() / 2 + 64);
    z.push_back(0x78); z.push_back(0x01);
    BitWriter w(z);
    w.put(1, 1); w.put(1, 2);                                     /* final block, fixed Huffman codes */
    const size_t dists[2] = {4, stride <= 32768 ? stride : 4};
    for (size_t i = 0; i < raw.size();) {
        size_t bestLen = 0, bestDist = 0;
        for (size_t d : dists) {
            if (d > i) continue;
            size_t len = 0;
            const size_t cap = std::min<size_t>(258, raw.size() - i);
            while (len < cap && raw[i + len] == raw[i + len - d]) len++;
            if (len > bestLen) { bestLen = len; bestDist = d; }
        }
        if (bestLen >= 4) { putLength(w, (uint32_t)bestLen); putDistance(w, (uint32_t)bestDist); i += bestLen; }
        else putLiteral(w, raw[i++]);
    }
    putLiteral(w, 256);                                           /* end of block */
    w.flush();
    const uint32_t adler = (b << 16) | a;
    for (int s = 24; s >= 0; s -= 8) z.push_back((uint8_t)(adler >> s));

    std::vector<uint8_t> file = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr(13);
    for (int k = 0; k < 4; k++) { ihdr[k] = (uint8_t)(width >> (24 - 8 * k)); ihdr[4 + k] = (uint8_t)(height >> (24 - 8 * k)); }
    ihdr[8] = 8; ihdr[9] = 6; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0;      /* 8 bits, RGBA, deflate, adaptive filters, no interlace */
    chunk(file, "IHDR", ihdr);
    chunk(file, "IDAT", z);
    chunk(file, "IEND", std::vector<uint8_t>());
    FILE* f = std::fopen(path, "wb");
    if (!f) { setError(std::string("okrt_write_png: cannot open ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    const bool ok = std::fwrite(file.data(), 1, file.size(), f) == file.size();
    std::fclose(f);
    if (!ok) { setError(std::string("okrt_write_png: short write to ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    return OKRT_OK;
}

/* Progressive state on disk: "OKRTACC1", width, height, numAccumulationFrames (u32 each), then W*H RGBA32F (u1).  */
extern "C" OKRT_API int okrt_write_accumulation_file(const char* path, const float* rgba32f, uint32_t width, uint32_t height, uint32_t numAccumulationFrames)
{
    using okrt::setError;
    if (!path || !rgba32f || !width || !height) { setError("okrt_write_accumulation_file: null or empty argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { setError(std::string("okrt_write_accumulation_file: cannot open ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    const uint32_t head[3] = {width, height, numAccumulationFrames};
    const size_t n = (size_t)width * height * 4;
    const bool ok = std::fwrite("OKRTACC1", 1, 8, f) == 8 && std::fwrite(head, 4, 3, f) == 3 && std::fwrite(rgba32f, 4, n, f) == n;
    std::fclose(f);
    if (!ok) { setError(std::string("okrt_write_accumulation_file: short write to ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    return OKRT_OK;
}

/* rgba32f == NULL: only the header is read (size the buffer, then call again) */
extern "C" OKRT_API int okrt_read_accumulation_file(const char* path, float* rgba32f, uint32_t* width, uint32_t* height, uint32_t* numAccumulationFrames)
{
    using okrt::setError;
    if (!path || !width || !height || !numAccumulationFrames) { setError("okrt_read_accumulation_file: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    FILE* f = std::fopen(path, "rb");
    if (!f) { setError(std::string("okrt_read_accumulation_file: cannot open ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    char magic[8]; uint32_t head[3];
    bool ok = std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "OKRTACC1", 8) == 0 && std::fread(head, 4, 3, f) == 3 && head[0] && head[1];
    if (ok && rgba32f) {
        if (head[0] != *width || head[1] != *height) { std::fclose(f); setError("okrt_read_accumulation_file: buffer size differs from the file's"); return OKRT_ERR_INVALID_ARGUMENT; }
        const size_t n = (size_t)head[0] * head[1] * 4;
        ok = std::fread(rgba32f, 4, n, f) == n;
    }
    std::fclose(f);
    if (!ok) { setError(std::string("okrt_read_accumulation_file: not a complete accumulation file: ") + path); return OKRT_ERR_INVALID_ARGUMENT; }
    *width = head[0]; *height = head[1]; *numAccumulationFrames = head[2];
    return OKRT_OK;
}
