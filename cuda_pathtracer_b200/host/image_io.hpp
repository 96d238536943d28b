// image_io.hpp -- the output side of the render loop: float SUM image -> 8-bit PNG / Radiance HDR.
//
// Mirrors `class Image` and `saveImage()` of the reference (path_tracer/src/image.h:8-19,
// image.cpp:22-49, main.cpp:88-112): pixel (x, y) of the float SUM image divided by the sample
// count lands at (W-1-x, y) (the reference mirrors x), is clamped to [0, 1], multiplied by 255.f
// and truncated to a byte.  The PNG container is written here without any third-party code
// (zlib "stored" blocks: the file is larger than stb_image_write's, the decoded pixels are the same).
#pragma once
#include <string>
#include <vector>

namespace b200host {

class Image {
public:
    Image(int x, int y);
    void setPixel(int x, int y, float r, float g, float b);
    void savePNG(const std::string& baseFilename) const;     // writes baseFilename + ".png"
    void saveHDR(const std::string& baseFilename) const;     // writes baseFilename + ".hdr" (flat RGBE)
    // the 8-bit RGB bytes savePNG encodes (row-major, 3 bytes per pixel)
    std::vector<unsigned char> bytes() const;

private:
    int xSize, ySize;
    std::vector<float> pixels;
};

// saveImage (main.cpp:88-112): mean = sum / samples, x mirrored
Image image_from_sum(const float* sum_rgb, int width, int height, float samples);

// "<imageName>.<startTimeString>.<samples>samp" (main.cpp:104-108)
std::string output_basename(const std::string& imageName, const std::string& startTimeString, float samples);

// "%Y-%m-%d_%H-%M-%Sz" in UTC (preview.cpp:21-28)
std::string currentTimeString();

}  // namespace b200host
