// image_io.h -- film file writers that need no device (PNG); PPM / PFM live in api.cu next to the film read-back.
#pragma once
#include <cstdint>
#include <vector>

namespace ptx {

// 8-bit RGB PNG of a tightly packed w*h*3 image.  flip_y writes the rows bottom-up (what the reference's
// stbi_flip_vertically_on_write(1) does for its bottom-row-first h_image, ui_windows.h:205).
void encode_png_rgb8(int w, int h, const uint8_t* rgb, bool flip_y, std::vector<uint8_t>& out);
bool write_png_rgb8(const char* path, int w, int h, const uint8_t* rgb, bool flip_y);

}  // namespace ptx
