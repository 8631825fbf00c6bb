// image_decode.h — stb_image.load_from_memory_rgba (src/stb_image.zig:11-18) restated: PNG and
// baseline/progressive-free JPEG to tightly packed RGBA8, row 0 first.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace pt {
bool decode_image_rgba8(const uint8_t* data, size_t size, std::vector<uint8_t>& rgba, uint32_t& width, uint32_t& height,
                        std::string& error);
}
