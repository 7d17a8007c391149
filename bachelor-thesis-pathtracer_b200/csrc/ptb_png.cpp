// ptb_png.cpp — PNG files to 8-bit RGBA for the scene loader's textures (see ptb_png.h).
// Stands in for stbi_load(path, &w, &h, &channels, STBI_rgb_alpha) as loadTexture calls it (reference:
// src/types/texture.cpp:33-35; stb_image is a vendored third-party header there).  Restated from the PNG specification
// (ISO/IEC 15948) plus the three conventions of stb_image that decide the bytes: 16-bit samples keep their high byte,
// samples of 1/2/4 bits are scaled by 255 / (2^depth - 1), and a tRNS colour key is matched on the unscaled samples.
// zlib does the inflate.  Adam7-interlaced files are refused (none among the reference's assets).
#include "ptb_png.h"

#include <algorithm>
#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

// zlib is optional at build time: build.py probes for the header and the link library and defines PTB_NO_ZLIB when either is
// missing, so a box without the zlib development files still builds the renderer (PNG textures are then refused with a
// clear message; hand texels to ptb_upload_textures instead).
#if !defined(PTB_NO_ZLIB) && defined(__has_include)
#if !__has_include(<zlib.h>)
#define PTB_NO_ZLIB 1
#endif
#endif
#ifndef PTB_NO_ZLIB
#include <zlib.h>
#endif

namespace ptb {
namespace {

uint32_t be32(const unsigned char *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

} // namespace

bool decode_png(const std::string &path, std::vector<unsigned char> &rgba, int &width, int &height, std::string &err) {
#ifdef PTB_NO_ZLIB
    (void)rgba; (void)width; (void)height;
    err = path + ": PNG textures are unsupported in this build of libptb (compiled without zlib)";
    return false;
#else
    std::vector<unsigned char> file;
    {
        FILE *f = fopen(path.c_str(), "rb");
        if (!f) {
            err = "cannot open " + path;
            return false;
        }
        unsigned char buf[65536];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, f)) > 0) file.insert(file.end(), buf, buf + n);
        fclose(f);
    }
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n'};
    if (file.size() < 8 || memcmp(file.data(), sig, 8) != 0) {
        err = path + ": not a PNG file (PNG and TGA textures are decoded)";
        return false;
    }
    uint32_t w = 0, h = 0;
    int depth = 0, colour = 0, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    bool have_ihdr = false, done = false;
    for (size_t off = 8; !done && off + 12 <= file.size();) {
        const uint32_t len = be32(&file[off]);
        const unsigned char *tag = &file[off + 4], *body = &file[off + 8];
        if (off + 12 + (size_t)len > file.size()) {
            err = path + ": truncated PNG chunk";
            return false;
        }
        if (!memcmp(tag, "IHDR", 4) && len >= 13) {
            w = be32(body);
            h = be32(body + 4);
            depth = body[8];
            colour = body[9];
            interlace = body[12];
            have_ihdr = true;
        } else if (!memcmp(tag, "PLTE", 4)) {
            plte.assign(body, body + len);
        } else if (!memcmp(tag, "tRNS", 4)) {
            trns.assign(body, body + len);
        } else if (!memcmp(tag, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!memcmp(tag, "IEND", 4)) {
            done = true;
        }
        off += 12 + (size_t)len;
    }
    const int channels = colour == 0 ? 1 : colour == 2 ? 3 : colour == 3 ? 1 : colour == 4 ? 2 : colour == 6 ? 4 : 0;
    const bool depth_ok = colour == 3 ? (depth == 1 || depth == 2 || depth == 4 || depth == 8)
                                      : colour == 0 ? (depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16) : (depth == 8 || depth == 16);
    if (!have_ihdr || !channels || !depth_ok || w == 0 || h == 0 || w > 32768 || h > 32768) {
        err = path + ": unsupported or corrupt PNG header";
        return false;
    }
    if (interlace) {
        err = path + ": Adam7-interlaced PNG files are not decoded; re-save the texture without interlacing";
        return false;
    }
    if (colour == 3 && plte.size() < 3) {
        err = path + ": palette image without PLTE";
        return false;
    }
    const size_t bpp_bits = (size_t)channels * depth, stride = (w * bpp_bits + 7) / 8, bpp = (bpp_bits + 7) / 8;
    std::vector<unsigned char> raw((stride + 1) * (size_t)h);
    uLongf got = (uLongf)raw.size();
    const int z = uncompress(raw.data(), &got, idat.data(), (uLong)idat.size());
    if (z != Z_OK || got != raw.size()) {
        err = path + ": PNG pixel data does not inflate to the image size";
        return false;
    }
    // undo the per-row filters in place (rows start with their filter type)
    for (uint32_t y = 0; y < h; ++y) {
        unsigned char *row = &raw[(stride + 1) * (size_t)y + 1];
        const unsigned char *up = y ? row - (stride + 1) : nullptr;
        const int filter = row[-1];
        if (filter > 4) {
            err = path + ": bad PNG row filter";
            return false;
        }
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? row[i - bpp] : 0, b = up ? up[i] : 0, c = (up && i >= bpp) ? up[i - bpp] : 0;
            const int add = filter == 0 ? 0 : filter == 1 ? a : filter == 2 ? b : filter == 3 ? (a + b) / 2 : paeth(a, b, c);
            row[i] = (unsigned char)(row[i] + add);
        }
    }
    width = (int)w;
    height = (int)h;
    rgba.assign((size_t)w * h * 4, 255);
    const int maxv = (1 << depth) - 1;
    int key[3] = {-1, -1, -1}; // tRNS colour key of grey / RGB images, in sample units
    const int key_mask = depth == 16 ? 0xffff : 0xff; // stb keeps the low byte of the key for samples of up to 8 bits
    if (colour == 0 && trns.size() >= 2) key[0] = ((trns[0] << 8) | trns[1]) & key_mask;
    if (colour == 2 && trns.size() >= 6)
        for (int k = 0; k < 3; ++k) key[k] = ((trns[2 * k] << 8) | trns[2 * k + 1]) & key_mask;
    for (uint32_t y = 0; y < h; ++y) {
        const unsigned char *row = &raw[(stride + 1) * (size_t)y + 1];
        for (uint32_t x = 0; x < w; ++x) {
            int s[4] = {0, 0, 0, 0}; // the pixel's samples, full precision
            for (int k = 0; k < channels; ++k) {
                if (depth == 8) s[k] = row[(size_t)x * channels + k];
                else if (depth == 16) s[k] = (row[2 * ((size_t)x * channels + k)] << 8) | row[2 * ((size_t)x * channels + k) + 1];
                else {
                    const size_t bit = (size_t)x * depth;
                    s[k] = (row[bit >> 3] >> (8 - depth - (bit & 7))) & maxv;
                }
            }
            unsigned char *o = &rgba[((size_t)y * w + x) * 4];
            auto to8 = [&](int v) { return (unsigned char)(depth == 16 ? v >> 8 : depth == 8 ? v : v * (255 / maxv)); };
            switch (colour) {
            case 0:
                o[0] = o[1] = o[2] = to8(s[0]);
                if (key[0] >= 0 && s[0] == key[0]) o[3] = 0;
                break;
            case 2:
                for (int k = 0; k < 3; ++k) o[k] = to8(s[k]);
                if (key[0] >= 0 && s[0] == key[0] && s[1] == key[1] && s[2] == key[2]) o[3] = 0;
                break;
            case 3: {
                const size_t i = (size_t)s[0];
                if (3 * i + 2 < plte.size()) {
                    o[0] = plte[3 * i];
                    o[1] = plte[3 * i + 1];
                    o[2] = plte[3 * i + 2];
                } else {
                    o[0] = o[1] = o[2] = 0;
                }
                if (i < trns.size()) o[3] = trns[i];
                break;
            }
            case 4:
                o[0] = o[1] = o[2] = to8(s[0]);
                o[3] = to8(s[1]);
                break;
            default:
                for (int k = 0; k < 4; ++k) o[k] = to8(s[k]);
            }
        }
    }
    return true;
#endif
}

// Truevision TGA, the subset stb_image's loader (stbi__tga_load) is asked for by the reference's kind of assets: image types 2
// (true colour), 3 (grey) and their run-length flavours 10 / 11, at 24 / 32 / 8 bits per pixel; rows bottom-up unless bit 5
// of the descriptor says top-down; BGR(A) in the file.  Colour-mapped and 15/16-bit files are refused.
bool decode_tga(const std::string &path, std::vector<unsigned char> &rgba, int &width, int &height, std::string &err) {
    std::vector<unsigned char> file;
    {
        FILE *f = fopen(path.c_str(), "rb");
        if (!f) {
            err = "cannot open " + path;
            return false;
        }
        unsigned char buf[65536];
        size_t n;
        while ((n = fread(buf, 1, sizeof buf, f)) > 0) file.insert(file.end(), buf, buf + n);
        fclose(f);
    }
    if (file.size() < 18) {
        err = path + ": truncated TGA header";
        return false;
    }
    const int id_len = file[0], indexed = file[1];
    int type = file[2];
    const int w = file[12] | (file[13] << 8), h = file[14] | (file[15] << 8), bpp = file[16];
    const bool top_down = (file[17] >> 5) & 1;
    const bool rle = type >= 8;
    if (rle) type -= 8;
    const int comp = bpp == 8 ? 1 : bpp == 24 ? 3 : bpp == 32 ? 4 : 0;
    if (indexed || (type != 2 && type != 3) || !comp || w <= 0 || h <= 0) {
        err = path + ": unsupported TGA flavour (colour-mapped, 15/16-bit or empty)";
        return false;
    }
    const size_t npx = (size_t)w * h;
    std::vector<unsigned char> px(npx * comp);
    size_t off = 18 + (size_t)id_len;
    if (!rle) {
        if (off + px.size() > file.size()) {
            err = path + ": truncated TGA pixel data";
            return false;
        }
        memcpy(px.data(), &file[off], px.size());
    } else {
        for (size_t i = 0; i < npx;) { // packets: 1 + (cmd & 127) pixels, one value repeated if bit 7 is set
            if (off >= file.size()) {
                err = path + ": truncated TGA run-length data";
                return false;
            }
            const int cmd = file[off++];
            const size_t count = 1 + (size_t)(cmd & 127), n = std::min(count, npx - i);
            const size_t need = (cmd & 128) ? (size_t)comp : count * comp;
            if (off + need > file.size()) {
                err = path + ": truncated TGA run-length data";
                return false;
            }
            for (size_t k = 0; k < n; ++k) memcpy(&px[(i + k) * comp], &file[off + ((cmd & 128) ? 0 : k * comp)], (size_t)comp);
            off += need;
            i += n;
        }
    }
    width = w;
    height = h;
    rgba.assign(npx * 4, 255);
    for (int y = 0; y < h; ++y) {
        const unsigned char *row = &px[(size_t)(top_down ? y : h - 1 - y) * w * comp];
        for (int x = 0; x < w; ++x) {
            unsigned char *o = &rgba[((size_t)y * w + x) * 4];
            const unsigned char *q = row + (size_t)x * comp;
            if (comp == 1) {
                o[0] = o[1] = o[2] = q[0];
            } else {
                o[0] = q[2];
                o[1] = q[1];
                o[2] = q[0];
                if (comp == 4) o[3] = q[3];
            }
        }
    }
    return true;
}

bool decode_image(const std::string &path, std::vector<unsigned char> &rgba, int &width, int &height, std::string &err) {
    const size_t dot = path.find_last_of('.');
    std::string ext = dot == std::string::npos ? "" : path.substr(dot + 1);
    for (char &c : ext) c = (char)tolower((unsigned char)c);
    if (ext == "tga") return decode_tga(path, rgba, width, height, err); // TGA has no signature; stb tries it last, too
    return decode_png(path, rgba, width, height, err);
}

} // namespace ptb
