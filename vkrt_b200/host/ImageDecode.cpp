// ImageDecode.cpp — PNG (zlib) and JPEG (baseline + progressive Huffman, 8-bit) decoders for the glTF loader.
//
// The reference decodes textures through tinygltf -> stb_image v2.21 and forces 4 channels
// (submodules/tinygltf/tiny_gltf.h:2346-2444, req_comp = 4 at :2363).  This is an independent implementation of the
// published formats (ITU-T T.81, JFIF 1.02, RFC 2083/1950/1951 via zlib):
//   * inverse DCT: the Loeffler-Ligtenberg-Moschytz 11-multiply integer algorithm with 12-bit constants and the
//     pass-1/pass-2 rounding stb_image and IJG's "islow" use, so decoded texels agree with the reference's decoder;
//   * YCbCr -> RGB: JFIF equations in 20-bit fixed point;
//   * chroma upsampling (not used by the Sponza textures, which are all 4:4:4): sample replication.
// Images are decoded on worker threads by Model (row f-1 of SURVEY §8: parallel ingest).
#include "ImageDecode.hpp"

#include <zlib.h>

#include <algorithm>
#include <cstring>

namespace vkrt
{
namespace
{
// ------------------------------------------------------------------------------------------------------------------
// PNG
// ------------------------------------------------------------------------------------------------------------------
uint32_t be32(const uint8_t* p)
{
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

int paeth(int a, int b, int c)
{
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool decodePng(const uint8_t* data, size_t size, DecodedImage& out, std::string& err)
{
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (size < 8 || std::memcmp(data, sig, 8) != 0)
    {
        err = "not a PNG";
        return false;
    }
    uint32_t width = 0, height = 0;
    int depth = 0, colorType = 0, interlace = 0;
    std::vector<uint8_t> idat, palette, trns;
    size_t pos = 8;
    bool end = false;
    while (!end && pos + 12 <= size)
    {
        const uint32_t len = be32(data + pos);
        const uint8_t* type = data + pos + 4;
        const uint8_t* body = data + pos + 8;
        if (pos + 12 + (size_t)len > size)
        {
            err = "truncated PNG chunk";
            return false;
        }
        if (!std::memcmp(type, "IHDR", 4))
        {
            width = be32(body), height = be32(body + 4);
            depth = body[8], colorType = body[9], interlace = body[12];
        }
        else if (!std::memcmp(type, "PLTE", 4))
        {
            palette.assign(body, body + len);
        }
        else if (!std::memcmp(type, "tRNS", 4))
        {
            trns.assign(body, body + len);
        }
        else if (!std::memcmp(type, "IDAT", 4))
        {
            idat.insert(idat.end(), body, body + len);
        }
        else if (!std::memcmp(type, "IEND", 4))
        {
            end = true;
        }
        pos += 12 + (size_t)len;
    }
    if (width == 0 || height == 0 || interlace != 0 || (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16))
    {
        err = "unsupported PNG (interlaced or bad bit depth)";
        return false;
    }
    if (depth < 8 && colorType != 0 && colorType != 3)
    {
        err = "bad PNG depth for colour type";
        return false;
    }
    int channels = 0;
    switch (colorType)
    {
    case 0: channels = 1; break;
    case 2: channels = 3; break;
    case 3: channels = 1; break;
    case 4: channels = 2; break;
    case 6: channels = 4; break;
    default: err = "bad PNG colour type"; return false;
    }
    if (colorType == 3 && depth > 8)
    {
        err = "unsupported PNG palette depth";
        return false;
    }
    // filter unit is one byte for sub-byte depths; rows are packed MSB first
    const size_t bpp = std::max<size_t>(1, (size_t)channels * depth / 8), stride = ((size_t)channels * depth * width + 7) / 8;
    std::vector<uint8_t> raw((stride + 1) * (size_t)height);
    uLongf rawLen = (uLongf)raw.size();
    if (uncompress(raw.data(), &rawLen, idat.data(), (uLong)idat.size()) != Z_OK || rawLen != raw.size())
    {
        err = "PNG inflate failed";
        return false;
    }
    std::vector<uint8_t> prev(stride, 0), cur(stride);
    out.width = width, out.height = height;
    out.rgba.assign((size_t)width * height * 4, 255);
    for (uint32_t y = 0; y < height; ++y)
    {
        const uint8_t* row = raw.data() + (stride + 1) * (size_t)y;
        const int filter = row[0];
        for (size_t i = 0; i < stride; ++i)
        {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int v = row[1 + i];
            switch (filter)
            {
            case 1: v += a; break;
            case 2: v += b; break;
            case 3: v += (a + b) >> 1; break;
            case 4: v += paeth(a, b, c); break;
            default: break;
            }
            cur[i] = (uint8_t)v;
        }
        uint8_t* dst = out.rgba.data() + (size_t)y * width * 4;
        const size_t step = depth / 8; // 16-bit: keep the high byte
        for (uint32_t x = 0; x < width; ++x)
        {
            uint8_t unpacked[1];
            const uint8_t* s = cur.data() + (size_t)x * bpp;
            if (depth < 8)
            {
                const size_t bit = (size_t)x * depth;
                const int v = (cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1);
                unpacked[0] = (uint8_t)(colorType == 0 ? v * 255 / ((1 << depth) - 1) : v);
                s = unpacked;
            }
            switch (colorType)
            {
            case 0: dst[0] = dst[1] = dst[2] = s[0]; break;
            case 2: dst[0] = s[0], dst[1] = s[step], dst[2] = s[2 * step]; break;
            case 3:
                if ((size_t)s[0] * 3 + 2 < palette.size())
                {
                    dst[0] = palette[s[0] * 3], dst[1] = palette[s[0] * 3 + 1], dst[2] = palette[s[0] * 3 + 2];
                }
                if (s[0] < trns.size())
                {
                    dst[3] = trns[s[0]];
                }
                break;
            case 4: dst[0] = dst[1] = dst[2] = s[0], dst[3] = s[step]; break;
            case 6: dst[0] = s[0], dst[1] = s[step], dst[2] = s[2 * step], dst[3] = s[3 * step]; break;
            }
            dst += 4;
        }
        prev.swap(cur);
    }
    return true;
}

// ------------------------------------------------------------------------------------------------------------------
// JPEG
// ------------------------------------------------------------------------------------------------------------------
const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                             41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                             30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct Huffman
{
    // canonical decode tables of T.81 F.2.2.3
    int mincode[17], maxcode[18], valptr[17];
    uint8_t values[256];
    bool present = false;
    void build(const uint8_t counts[16], const uint8_t* symbols, int n)
    {
        std::memcpy(values, symbols, (size_t)n);
        int code = 0, k = 0;
        for (int len = 1; len <= 16; ++len)
        {
            valptr[len] = k;
            mincode[len] = code;
            code += counts[len - 1];
            k += counts[len - 1];
            maxcode[len] = counts[len - 1] ? code - 1 : -1;
            code <<= 1;
        }
        maxcode[17] = 0x7FFFFFFF;
        present = true;
    }
};

struct BitReader
{
    const uint8_t* p;
    const uint8_t* end;
    uint32_t acc = 0;
    int bits = 0;
    bool hitMarker = false;
    int marker = 0;
    void fill()
    {
        while (bits <= 24)
        {
            int byte = 0;
            if (!hitMarker && p < end)
            {
                byte = *p++;
                if (byte == 0xFF)
                {
                    int next = p < end ? *p : 0xD9;
                    if (next == 0)
                    {
                        ++p; // stuffed zero
                    }
                    else
                    {
                        hitMarker = true; // leave the marker for the segment parser; feed zeros from here on
                        marker = next;
                        --p;
                        byte = 0;
                    }
                }
            }
            acc |= (uint32_t)byte << (24 - bits);
            bits += 8;
        }
    }
    int bit()
    {
        if (bits < 1)
        {
            fill();
        }
        const int b = (int)(acc >> 31);
        acc <<= 1;
        --bits;
        return b;
    }
    int get(int n)
    {
        if (n == 0)
        {
            return 0;
        }
        if (bits < n)
        {
            fill();
        }
        const int v = (int)(acc >> (32 - n));
        acc <<= n;
        bits -= n;
        return v;
    }
    void reset()
    {
        acc = 0;
        bits = 0;
        hitMarker = false;
    }
};

inline int extend(int v, int n) // T.81 F.2.2.1 EXTEND
{
    return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
}

int decodeSymbol(BitReader& br, const Huffman& h)
{
    int code = 0;
    for (int len = 1; len <= 16; ++len)
    {
        code = (code << 1) | br.bit();
        if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len])
        {
            return h.values[h.valptr[len] + code - h.mincode[len]];
        }
    }
    return -1;
}

struct Component
{
    int id = 0, h = 1, v = 1, tq = 0;
    int td = 0, ta = 0;      // tables selected by the current scan
    int blocksW = 0, blocksH = 0; // allocated blocks (padded to whole MCUs)
    int pred = 0;
    std::vector<int16_t> coef; // blocksW*blocksH*64, natural (de-zigzagged) order
    std::vector<uint8_t> plane; // blocksW*8 x blocksH*8 samples
};

// LL&M integer inverse DCT, 12-bit constants; column pass keeps 2 extra bits, row pass adds the +128 level shift.
inline int f2f(double x) { return (int)(x * 4096 + 0.5); }

struct IdctConstants
{
    int c0541, c1847, c0765, c1175, c0298, c2053, c3072, c1501, c0899, c2562, c1961, c0390;
    IdctConstants()
    {
        c0541 = f2f(0.5411961), c1847 = f2f(-1.847759065), c0765 = f2f(0.765366865), c1175 = f2f(1.175875602);
        c0298 = f2f(0.298631336), c2053 = f2f(2.053119869), c3072 = f2f(3.072711026), c1501 = f2f(1.501321110);
        c0899 = f2f(-0.899976223), c2562 = f2f(-2.562915447), c1961 = f2f(-1.961570560), c0390 = f2f(-0.390180644);
    }
};
const IdctConstants K;

inline void idct1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7, int& x0, int& x1, int& x2, int& x3, int& t0, int& t1,
                   int& t2, int& t3)
{
    int p2 = s2, p3 = s6;
    int p1 = (p2 + p3) * K.c0541;
    t2 = p1 + p3 * K.c1847;
    t3 = p1 + p2 * K.c0765;
    p2 = s0, p3 = s4;
    t0 = (p2 + p3) * 4096;
    t1 = (p2 - p3) * 4096;
    x0 = t0 + t3, x3 = t0 - t3, x1 = t1 + t2, x2 = t1 - t2;
    t0 = s7, t1 = s5, t2 = s3, t3 = s1;
    p3 = t0 + t2;
    int p4 = t1 + t3;
    p1 = t0 + t3;
    p2 = t1 + t2;
    const int p5 = (p3 + p4) * K.c1175;
    t0 = t0 * K.c0298, t1 = t1 * K.c2053, t2 = t2 * K.c3072, t3 = t3 * K.c1501;
    p1 = p5 + p1 * K.c0899;
    p2 = p5 + p2 * K.c2562;
    p3 = p3 * K.c1961;
    p4 = p4 * K.c0390;
    t3 += p1 + p4, t2 += p2 + p3, t1 += p2 + p4, t0 += p1 + p3;
}

inline uint8_t clamp8(int x)
{
    return (uint8_t)(x < 0 ? 0 : (x > 255 ? 255 : x));
}

void idctBlock(uint8_t* out, int outStride, const int16_t* coef, const uint16_t* quant)
{
    int val[64];
    int d[64];
    for (int i = 0; i < 64; ++i)
    {
        d[i] = coef[i] * quant[i];
    }
    for (int c = 0; c < 8; ++c) // columns
    {
        const int* s = d + c;
        int* v = val + c;
        if (s[8] == 0 && s[16] == 0 && s[24] == 0 && s[32] == 0 && s[40] == 0 && s[48] == 0 && s[56] == 0)
        {
            const int dc = s[0] * 4;
            v[0] = v[8] = v[16] = v[24] = v[32] = v[40] = v[48] = v[56] = dc;
            continue;
        }
        int x0, x1, x2, x3, t0, t1, t2, t3;
        idct1d(s[0], s[8], s[16], s[24], s[32], s[40], s[48], s[56], x0, x1, x2, x3, t0, t1, t2, t3);
        x0 += 512, x1 += 512, x2 += 512, x3 += 512;
        v[0] = (x0 + t3) >> 10, v[56] = (x0 - t3) >> 10;
        v[8] = (x1 + t2) >> 10, v[48] = (x1 - t2) >> 10;
        v[16] = (x2 + t1) >> 10, v[40] = (x2 - t1) >> 10;
        v[24] = (x3 + t0) >> 10, v[32] = (x3 - t0) >> 10;
    }
    for (int r = 0; r < 8; ++r) // rows
    {
        const int* v = val + r * 8;
        uint8_t* o = out + r * outStride;
        int x0, x1, x2, x3, t0, t1, t2, t3;
        idct1d(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], x0, x1, x2, x3, t0, t1, t2, t3);
        const int bias = 65536 + (128 << 17);
        x0 += bias, x1 += bias, x2 += bias, x3 += bias;
        o[0] = clamp8((x0 + t3) >> 17), o[7] = clamp8((x0 - t3) >> 17);
        o[1] = clamp8((x1 + t2) >> 17), o[6] = clamp8((x1 - t2) >> 17);
        o[2] = clamp8((x2 + t1) >> 17), o[5] = clamp8((x2 - t1) >> 17);
        o[3] = clamp8((x3 + t0) >> 17), o[4] = clamp8((x3 - t0) >> 17);
    }
}

struct Jpeg
{
    int width = 0, height = 0;
    bool progressive = false;
    int ncomp = 0;
    Component comp[4];
    uint16_t quant[4][64];
    Huffman dc[4], ac[4];
    int hmax = 1, vmax = 1, mcuW = 0, mcuH = 0, mcusX = 0, mcusY = 0;
    int restartInterval = 0;
    int eobrun = 0;
    // current scan
    int scanN = 0, scanOrder[4];
    int ss = 0, se = 63, ah = 0, al = 0;
};

bool decodeBlockBaseline(Jpeg& j, BitReader& br, Component& c, int16_t* blk)
{
    const int t = decodeSymbol(br, j.dc[c.td]);
    if (t < 0 || t > 15)
    {
        return false;
    }
    const int diff = t ? extend(br.get(t), t) : 0;
    c.pred += diff;
    blk[0] = (int16_t)c.pred;
    for (int k = 1; k < 64;)
    {
        const int rs = decodeSymbol(br, j.ac[c.ta]);
        if (rs < 0)
        {
            return false;
        }
        const int r = rs >> 4, s = rs & 15;
        if (s == 0)
        {
            if (r != 15)
            {
                break; // EOB
            }
            k += 16;
            continue;
        }
        k += r;
        if (k > 63)
        {
            return false;
        }
        blk[kZigzag[k++]] = (int16_t)extend(br.get(s), s);
    }
    return true;
}

bool decodeBlockProgDc(Jpeg& j, BitReader& br, Component& c, int16_t* blk)
{
    if (j.ah == 0)
    {
        const int t = decodeSymbol(br, j.dc[c.td]);
        if (t < 0 || t > 15)
        {
            return false;
        }
        const int diff = t ? extend(br.get(t), t) : 0;
        c.pred += diff;
        blk[0] = (int16_t)(c.pred * (1 << j.al));
    }
    else if (br.bit())
    {
        blk[0] = (int16_t)(blk[0] + (1 << j.al));
    }
    return true;
}

bool decodeBlockProgAc(Jpeg& j, BitReader& br, Component& c, int16_t* blk)
{
    if (j.ah == 0) // first pass of this band
    {
        if (j.eobrun)
        {
            --j.eobrun;
            return true;
        }
        for (int k = j.ss; k <= j.se;)
        {
            const int rs = decodeSymbol(br, j.ac[c.ta]);
            if (rs < 0)
            {
                return false;
            }
            const int r = rs >> 4, s = rs & 15;
            if (s == 0)
            {
                if (r < 15)
                {
                    j.eobrun = (1 << r) - 1;
                    if (r)
                    {
                        j.eobrun += br.get(r);
                    }
                    break;
                }
                k += 16;
            }
            else
            {
                k += r;
                if (k > 63)
                {
                    return false;
                }
                blk[kZigzag[k++]] = (int16_t)(extend(br.get(s), s) * (1 << j.al));
            }
        }
        return true;
    }
    // refinement pass (T.81 G.1.2.3)
    const int bitv = 1 << j.al;
    int k = j.ss;
    if (j.eobrun == 0)
    {
        for (; k <= j.se;)
        {
            const int rs = decodeSymbol(br, j.ac[c.ta]);
            if (rs < 0)
            {
                return false;
            }
            int r = rs >> 4;
            const int s = rs & 15;
            int value = 0;
            if (s == 0)
            {
                if (r < 15)
                {
                    j.eobrun = (1 << r);
                    if (r)
                    {
                        j.eobrun += br.get(r);
                    }
                    break; // the rest of this block is handled by the EOB-run branch below
                }
            }
            else
            {
                if (s != 1)
                {
                    return false;
                }
                value = br.bit() ? bitv : -bitv;
            }
            // skip r zero-history coefficients, refining the non-zero ones on the way
            while (k <= j.se)
            {
                int16_t& p = blk[kZigzag[k++]];
                if (p != 0)
                {
                    if (br.bit() && (p & bitv) == 0)
                    {
                        p = (int16_t)(p > 0 ? p + bitv : p - bitv);
                    }
                }
                else
                {
                    if (r == 0)
                    {
                        if (value)
                        {
                            p = (int16_t)value;
                        }
                        break;
                    }
                    --r;
                }
            }
        }
    }
    if (j.eobrun)
    {
        for (; k <= j.se; ++k)
        {
            int16_t& p = blk[kZigzag[k]];
            if (p != 0 && br.bit() && (p & bitv) == 0)
            {
                p = (int16_t)(p > 0 ? p + bitv : p - bitv);
            }
        }
        --j.eobrun;
    }
    return true;
}

bool decodeScan(Jpeg& j, BitReader& br)
{
    for (int i = 0; i < j.scanN; ++i)
    {
        j.comp[j.scanOrder[i]].pred = 0;
    }
    j.eobrun = 0;
    int restartCount = j.restartInterval;
    auto restart = [&]() -> bool {
        if (j.restartInterval && --restartCount == 0)
        {
            // align to a byte boundary and consume the RSTn marker
            br.bits = 0, br.acc = 0;
            if (!br.hitMarker)
            {
                // look ahead for the marker in the byte stream
                while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7))
                {
                    ++br.p;
                }
            }
            if (br.p + 1 < br.end && br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)
            {
                br.p += 2;
            }
            br.reset();
            for (int i = 0; i < j.scanN; ++i)
            {
                j.comp[j.scanOrder[i]].pred = 0;
            }
            j.eobrun = 0;
            restartCount = j.restartInterval;
        }
        return true;
    };
    auto decodeOne = [&](Component& c, int bx, int by) -> bool {
        int16_t* blk = c.coef.data() + ((size_t)by * c.blocksW + bx) * 64;
        if (!j.progressive)
        {
            return decodeBlockBaseline(j, br, c, blk);
        }
        return j.ss == 0 ? decodeBlockProgDc(j, br, c, blk) : decodeBlockProgAc(j, br, c, blk);
    };
    if (j.scanN == 1)
    {
        // non-interleaved: the component's own blocks in raster order, only those covering the image
        Component& c = j.comp[j.scanOrder[0]];
        const int w = (((j.width * c.h + j.hmax - 1) / j.hmax) + 7) >> 3;
        const int h = (((j.height * c.v + j.vmax - 1) / j.vmax) + 7) >> 3;
        for (int by = 0; by < h; ++by)
        {
            for (int bx = 0; bx < w; ++bx)
            {
                if (!decodeOne(c, bx, by))
                {
                    return false;
                }
                restart();
            }
        }
        return true;
    }
    for (int my = 0; my < j.mcusY; ++my)
    {
        for (int mx = 0; mx < j.mcusX; ++mx)
        {
            for (int i = 0; i < j.scanN; ++i)
            {
                Component& c = j.comp[j.scanOrder[i]];
                for (int y = 0; y < c.v; ++y)
                {
                    for (int x = 0; x < c.h; ++x)
                    {
                        if (!decodeOne(c, mx * c.h + x, my * c.v + y))
                        {
                            return false;
                        }
                    }
                }
            }
            restart();
        }
    }
    return true;
}

bool decodeJpeg(const uint8_t* data, size_t size, DecodedImage& out, std::string& err)
{
    if (size < 4 || data[0] != 0xFF || data[1] != 0xD8)
    {
        err = "not a JPEG";
        return false;
    }
    Jpeg j;
    std::memset(j.quant, 0, sizeof(j.quant));
    size_t pos = 2;
    bool haveFrame = false;
    while (pos + 4 <= size)
    {
        if (data[pos] != 0xFF)
        {
            ++pos;
            continue;
        }
        const int m = data[pos + 1];
        if (m == 0xFF || m == 0x00 || (m >= 0xD0 && m <= 0xD7) || m == 0x01)
        {
            pos += (m == 0xFF) ? 1 : 2;
            continue;
        }
        if (m == 0xD9)
        {
            break;
        }
        const size_t len = ((size_t)data[pos + 2] << 8) | data[pos + 3];
        const uint8_t* seg = data + pos + 4;
        const uint8_t* segEnd = data + pos + 2 + len;
        if (pos + 2 + len > size || len < 2)
        {
            err = "truncated JPEG segment";
            return false;
        }
        if (m == 0xDB) // DQT
        {
            while (seg < segEnd)
            {
                const int pq = seg[0] >> 4, tq = seg[0] & 15;
                ++seg;
                if (tq > 3)
                {
                    err = "bad DQT";
                    return false;
                }
                for (int i = 0; i < 64; ++i)
                {
                    j.quant[tq][kZigzag[i]] = pq ? (uint16_t)((seg[0] << 8) | seg[1]) : seg[0];
                    seg += pq ? 2 : 1;
                }
            }
        }
        else if (m == 0xC4) // DHT
        {
            while (seg + 17 <= segEnd)
            {
                const int tc = seg[0] >> 4, th = seg[0] & 15;
                int n = 0;
                for (int i = 0; i < 16; ++i)
                {
                    n += seg[1 + i];
                }
                if (th > 3 || n > 256 || seg + 17 + n > segEnd)
                {
                    err = "bad DHT";
                    return false;
                }
                (tc ? j.ac[th] : j.dc[th]).build(seg + 1, seg + 17, n);
                seg += 17 + n;
            }
        }
        else if (m == 0xC0 || m == 0xC1 || m == 0xC2) // SOF0/1/2
        {
            if (seg[0] != 8)
            {
                err = "only 8-bit JPEG supported";
                return false;
            }
            j.progressive = (m == 0xC2);
            j.height = (seg[1] << 8) | seg[2];
            j.width = (seg[3] << 8) | seg[4];
            j.ncomp = seg[5];
            if (j.width <= 0 || j.height <= 0 || (j.ncomp != 1 && j.ncomp != 3))
            {
                err = "unsupported JPEG frame";
                return false;
            }
            for (int i = 0; i < j.ncomp; ++i)
            {
                Component& c = j.comp[i];
                c.id = seg[6 + 3 * i];
                c.h = seg[7 + 3 * i] >> 4;
                c.v = seg[7 + 3 * i] & 15;
                c.tq = seg[8 + 3 * i] & 3;
                if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4)
                {
                    err = "bad sampling factors";
                    return false;
                }
                j.hmax = std::max(j.hmax, c.h), j.vmax = std::max(j.vmax, c.v);
            }
            j.mcuW = 8 * j.hmax, j.mcuH = 8 * j.vmax;
            j.mcusX = (j.width + j.mcuW - 1) / j.mcuW, j.mcusY = (j.height + j.mcuH - 1) / j.mcuH;
            for (int i = 0; i < j.ncomp; ++i)
            {
                Component& c = j.comp[i];
                c.blocksW = j.mcusX * c.h, c.blocksH = j.mcusY * c.v;
                c.coef.assign((size_t)c.blocksW * c.blocksH * 64, 0);
            }
            haveFrame = true;
        }
        else if (m == 0xC3 || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC))
        {
            err = "unsupported JPEG process (lossless / hierarchical / arithmetic)";
            return false;
        }
        else if (m == 0xDD) // DRI
        {
            j.restartInterval = (seg[0] << 8) | seg[1];
        }
        else if (m == 0xDA) // SOS
        {
            if (!haveFrame)
            {
                err = "SOS before SOF";
                return false;
            }
            j.scanN = seg[0];
            if (j.scanN < 1 || j.scanN > j.ncomp)
            {
                err = "bad SOS";
                return false;
            }
            for (int i = 0; i < j.scanN; ++i)
            {
                const int id = seg[1 + 2 * i];
                int which = -1;
                for (int k = 0; k < j.ncomp; ++k)
                {
                    if (j.comp[k].id == id)
                    {
                        which = k;
                    }
                }
                if (which < 0)
                {
                    err = "SOS names an unknown component";
                    return false;
                }
                j.scanOrder[i] = which;
                j.comp[which].td = seg[2 + 2 * i] >> 4;
                j.comp[which].ta = seg[2 + 2 * i] & 15;
                if (j.comp[which].td > 3 || j.comp[which].ta > 3)
                {
                    err = "bad table selector";
                    return false;
                }
            }
            j.ss = seg[1 + 2 * j.scanN];
            j.se = seg[2 + 2 * j.scanN];
            j.ah = seg[3 + 2 * j.scanN] >> 4;
            j.al = seg[3 + 2 * j.scanN] & 15;
            if (!j.progressive)
            {
                j.ss = 0, j.se = 63, j.ah = 0, j.al = 0;
            }
            BitReader br;
            br.p = segEnd;
            br.end = data + size;
            if (!decodeScan(j, br))
            {
                err = "corrupt JPEG entropy data";
                return false;
            }
            // continue parsing at the marker that ended the scan
            size_t q = (size_t)(br.p - data);
            while (q + 1 < size && !(data[q] == 0xFF && data[q + 1] != 0x00 && !(data[q + 1] >= 0xD0 && data[q + 1] <= 0xD7)))
            {
                ++q;
            }
            pos = q;
            continue;
        }
        pos += 2 + len;
    }
    if (!haveFrame)
    {
        err = "JPEG without a frame header";
        return false;
    }
    // dequantise + inverse DCT into per-component planes
    for (int i = 0; i < j.ncomp; ++i)
    {
        Component& c = j.comp[i];
        const int pw = c.blocksW * 8;
        c.plane.assign((size_t)pw * c.blocksH * 8, 0);
        for (int by = 0; by < c.blocksH; ++by)
        {
            for (int bx = 0; bx < c.blocksW; ++bx)
            {
                idctBlock(c.plane.data() + ((size_t)by * 8) * pw + (size_t)bx * 8, pw, c.coef.data() + ((size_t)by * c.blocksW + bx) * 64,
                          j.quant[c.tq]);
            }
        }
    }
    out.width = (uint32_t)j.width, out.height = (uint32_t)j.height;
    out.rgba.resize((size_t)j.width * j.height * 4);
    auto sample = [&](const Component& c, int x, int y) -> int {
        const int sx = x * c.h / j.hmax, sy = y * c.v / j.vmax; // replication upsampling
        return c.plane[(size_t)sy * (c.blocksW * 8) + sx];
    };
    // JFIF YCbCr -> RGB in 20-bit fixed point
    const int crR = ((int)(1.40200f * 4096.0f + 0.5f)) << 8, crG = -(((int)(0.71414f * 4096.0f + 0.5f)) << 8);
    const int cbG = -(((int)(0.34414f * 4096.0f + 0.5f)) << 8), cbB = ((int)(1.77200f * 4096.0f + 0.5f)) << 8;
    for (int y = 0; y < j.height; ++y)
    {
        uint8_t* dst = out.rgba.data() + (size_t)y * j.width * 4;
        for (int x = 0; x < j.width; ++x)
        {
            if (j.ncomp == 1)
            {
                const uint8_t g = (uint8_t)sample(j.comp[0], x, y);
                dst[0] = dst[1] = dst[2] = g;
            }
            else
            {
                const int yy = (sample(j.comp[0], x, y) << 20) + (1 << 19);
                const int cb = sample(j.comp[1], x, y) - 128, cr = sample(j.comp[2], x, y) - 128;
                const int r = (yy + cr * crR) >> 20;
                const int g = (yy + cr * crG + (int)(((unsigned)(cb * cbG)) & 0xFFFF0000u)) >> 20;
                const int b = (yy + cb * cbB) >> 20;
                dst[0] = clamp8(r), dst[1] = clamp8(g), dst[2] = clamp8(b);
            }
            dst[3] = 255;
            dst += 4;
        }
    }
    return true;
}
} // namespace

bool decodeImage(const uint8_t* data, size_t size, DecodedImage& out, std::string& err)
{
    if (size >= 8 && data[0] == 0x89 && data[1] == 'P')
    {
        return decodePng(data, size, out, err);
    }
    if (size >= 3 && data[0] == 0xFF && data[1] == 0xD8)
    {
        return decodeJpeg(data, size, out, err);
    }
    err = "unknown image format";
    return false;
}

} // namespace vkrt
