// See ref_glue.hpp.  Replacements for reference src/progressview.cpp (nanogui) and src/texture.cpp (DevIL).
#include "ref_glue.hpp"
#include <cstring>

#define STB_IMAGE_IMPLEMENTATION
#define STBI_ONLY_JPEG
#define STBI_ONLY_PNG
#define STBI_ONLY_BMP
#define STBI_ONLY_TGA
#include "stb_image.h" // vendored inside the reference tree (ext/nanogui/ext/nanovg/src)

ProgressView::ProgressView(nanogui::Screen *) {}
void ProgressView::showError(const std::string &) {}
void ProgressView::showMessage(const std::string &, float) {}
void ProgressView::showMessage(const std::string &, const std::string &) {}
void ProgressView::showMessage(const std::string &, const std::string &, float) {}
void ProgressView::center() {}
void ProgressView::hide() {}

// Same contract as reference src/texture.cpp:16-41: RGBA8, top-left origin (DevIL's default), name
// "error" and a null data pointer when the file cannot be read.
Texture::Texture(const std::string path, const std::string filename)
{
    int w = 0, h = 0, n = 0;
    unsigned char *px = stbi_load(path.c_str(), &w, &h, &n, 4);
    if (px)
    {
        width = (cl_uint)w; height = (cl_uint)h;
        data = new cl_uchar[(size_t)w * h * 4];
        memcpy(data, px, (size_t)w * h * 4);
        stbi_image_free(px);
        name = filename;
    }
    else
    {
        width = height = 0; data = nullptr; name = "error";
    }
}
