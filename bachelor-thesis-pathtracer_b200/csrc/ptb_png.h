// ptb_png.h — PNG files to 8-bit RGBA, rows top to bottom: what stbi_load(path, .., STBI_rgb_alpha) hands loadTexture
// (reference: src/types/texture.cpp:33-35) for PNG input.  All colour types, 1-16 bits per sample, PLTE / tRNS;
// no Adam7.  Host-only; zlib for the inflate.
#pragma once
#include <string>
#include <vector>

namespace ptb {

bool decode_png(const std::string &path, std::vector<unsigned char> &rgba, int &width, int &height, std::string &err);
// true-colour / grey TGA (types 2, 3, 10, 11 at 24 / 32 / 8 bits), rows top to bottom, RGBA out
bool decode_tga(const std::string &path, std::vector<unsigned char> &rgba, int &width, int &height, std::string &err);
// by file name: .tga -> decode_tga, everything else must be PNG
bool decode_image(const std::string &path, std::vector<unsigned char> &rgba, int &width, int &height, std::string &err);

} // namespace ptb
