// ImageDecode.hpp — PNG / JPEG decode to RGBA8 for the glTF loader (see ImageDecode.cpp).
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace vkrt
{
struct DecodedImage
{
    uint32_t width = 0, height = 0;
    std::vector<uint8_t> rgba; // width*height*4, alpha 255 when the file has none
};

// Returns false and fills `err` for unsupported or corrupt files.
bool decodeImage(const uint8_t* data, size_t size, DecodedImage& out, std::string& err);
} // namespace vkrt
