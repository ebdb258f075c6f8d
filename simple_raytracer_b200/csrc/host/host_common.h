// host_common.h — shared bits of the CPU front-end (parser, P3 reader/writer).
#pragma once

#include "rt_host.h"

#include <string>
#include <vector>

namespace rthost {

// Owns every array an RtSceneDesc points to.  `desc` is the first member so the
// C ABI can hand out &desc and get the owner back with a cast.
struct OwnedScene {
    RtSceneDesc desc{};
    std::vector<RtMaterial> materials;
    std::vector<RtTexture> textures;
    std::vector<std::vector<uint8_t>> texture_pixels;
    std::vector<RtSphere> spheres;
    std::vector<RtFace> faces;
    std::vector<RtLight> lights;

    void seal();   // point desc at the vectors
};

void set_error(const std::string& msg);

// std::stof / std::stoi as the reference uses them: leading blanks skipped,
// trailing junk accepted ("1git" -> 1), no digits -> failure.
bool parse_float(const std::string& tok, float* out);
bool parse_int(const std::string& tok, int* out);

// utility.h:59-139 with hard errors where the reference has undefined behaviour.
int read_texture_p3(const std::string& path, int32_t* width, int32_t* height, std::vector<uint8_t>* rgb);

}  // namespace rthost
