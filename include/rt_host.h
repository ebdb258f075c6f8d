/*
 * rt_host.h — C ABI of the host front-end that stays on the CPU: the scene-file
 * parser, the P3 texture loader and the P3 writer of the reference, re-stated
 * with the same observable behaviour so that `raytracer scene.txt` keeps
 * working around the CUDA render core (rt.h).
 *
 *   scene parser     main.cpp:88-602        
 *   texture loader   src/utility.h:59-139   
 *   PPM writer       main.cpp:613-650       
 *
 * Nothing here touches the GPU; librt_host.so has no CUDA dependency.
 */
#ifndef RT_B200_RT_HOST_H
#define RT_B200_RT_HOST_H

#include "rt.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Parses a scene file into a freshly allocated RtSceneDesc (all arrays owned by
 * the returned object; free with rt_host_free_scene_desc).
 *
 * Behaviour follows the reference parser, including: split on single spaces,
 * an EMPTY token (double / leading space) is fatal (main.cpp:114), unknown
 * commands and commands without arguments are skipped (136), stof/stoi accept
 * trailing junk, mtlcolor alpha/eta only with exactly 12 arguments (293-299),
 * 1-based v/vn/vt maps where a missing index reads as zeros, per-vertex face
 * format sniffing (487-517), objects capture the current material/texture.
 * Where the reference aborts (uncaught exception) or prints and returns 0
 * without an image, this returns RT_E_PARSE with the reference's message in
 * rt_host_last_error().
 *
 * texture paths are resolved like the reference does: relative to the current
 * working directory — unless texture_dir is non-NULL, in which case relative
 * paths are resolved against it. */
int rt_host_parse_scene_file(const char* path, const char* texture_dir,
                             RtSceneDesc** out);
void rt_host_free_scene_desc(RtSceneDesc* desc);

/* P3 loader with the reference's acceptance rules (utility.h:59-139): "P3",
 * width, height, "255", then width*height*3 integers, tokens split on ' ' only,
 * lines starting with '#' skipped, an empty line is an error.  Differences by
 * design: a missing file, too few tokens or a value outside 0..255 are hard
 * errors here (undefined behaviour / out-of-range size_t in the reference).
 * rgb is malloc'ed: free with rt_host_free(). */
int rt_host_read_texture_p3(const char* path, int32_t* width, int32_t* height,
                            uint8_t** rgb);
void rt_host_free(void* p);

/* Writes the reference's P3 dialect byte for byte (main.cpp:628-641):
 * "P3 \n" "W H \n" "255 \n" then one "r g b \n" line per pixel.  Where
 * flags bit c (0..2) is set the token 18446744071562067968 is printed
 * ((size_t)(int)NaN on x86-64); where bit 3 is set the value is
 * (int)(bkg[c]*255.0f) so that an out-of-range bkgcolor prints as in the
 * reference.  flags may be NULL. */
int rt_host_write_ppm(const char* path, int32_t width, int32_t height,
                      const uint8_t* rgb, const uint8_t* flags,
                      const float bkg[3]);

/* Same text into a malloc'ed buffer (free with rt_host_free). */
int rt_host_format_ppm(int32_t width, int32_t height, const uint8_t* rgb,
                       const uint8_t* flags, const float bkg[3],
                       char** out_text, size_t* out_len);

/* "scene.txt" -> "scene.ppm" (utility.h:34-41, main.cpp:614-616). buf gets the
 * NUL-terminated result; returns RT_E_INVALID when it does not fit. */
int rt_host_output_path(const char* scene_path, char* buf, size_t buf_len);

const char* rt_host_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_RT_HOST_H */
