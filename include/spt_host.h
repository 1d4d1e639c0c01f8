/*
 * spt_host.h -- C exports of the host-side mirror (include/spt_host/*.hpp) for bindings and tools:
 * the reference's scene generators with an injectable seed, its camera basis, and its BMP output.
 *
 *   spt_host_init_spheres      InitSpheres()      /root/reference/include/SceneGenerators.hpp:68-133
 *   spt_host_generate_spheres  GenerateSpheres()  /root/reference/include/SceneGenerators.hpp:6-66
 *   spt_host_generate_metal_glass / _lattice       BASELINE.json configs 4 / 5 (no reference counterpart)
 *   spt_host_camera_basis      Transpose(CreateCameraBasisMatrix(...))  include/Math.hpp:198-231, Renderer.hpp:321
 *   spt_host_save_bmp          stbi_write_bmp as called by io::SaveImage  include/IOHelpers.hpp:24-27
 *   spt_host_save_scene / spt_host_load_scene   the SoA scene globals + camera as a file (include/spt_host/SceneFile.hpp)
 *   spt_host_save_pfm          linear-float companion of the BMP (no reference counterpart)
 *   spt_host_save_hdr          the same values as Radiance RGBE (what third_party/stb/stb_image_write.h:183 stbi_write_hdr serialises)
 *   spt_host_save_accum / spt_host_load_accum   accumulator checkpoint for resume (no reference counterpart)
 *
 * Generators return the sphere count; arrays are filled when cap >= count (call with cap = 0 to size).
 * Layout = the reference's SoA globals (include/Globals.hpp:31-37).
 */
#ifndef SPT_HOST_H
#define SPT_HOST_H

#include <stdint.h>

#ifndef SPT_API
#if defined(__GNUC__)
#define SPT_API __attribute__((visibility("default"))) /* the library is built with -fvisibility=hidden */
#else
#define SPT_API
#endif
#endif

#ifdef __cplusplus
extern "C" {
#endif

SPT_API uint32_t spt_host_init_spheres(uint32_t seed, uint32_t cap, float *centers4, float *radii, uint8_t *materials,
                               float *colors4, float *fuzz);
SPT_API uint32_t spt_host_generate_spheres(uint32_t seed, uint32_t cap, float *centers4, float *radii, uint8_t *materials,
                                   float *colors4, float *fuzz);
SPT_API uint32_t spt_host_generate_metal_glass(uint32_t seed, uint32_t count, uint32_t cap, float *centers4, float *radii,
                                       uint8_t *materials, float *colors4, float *fuzz);
SPT_API uint32_t spt_host_generate_lattice(uint32_t seed, uint32_t side, uint32_t cap, float *centers4, float *radii,
                                   uint8_t *materials, float *colors4, float *fuzz);
SPT_API void spt_host_camera_basis(const float eye[3], const float look[3], const float up[3], float view16[16]);
SPT_API int spt_host_save_bmp(const char *path, uint32_t width, uint32_t height, const uint8_t *rgb8);
/* scene file: camera16 = eye xyz0, lookAt xyz0, up xyz0, sky rgb0.  load: call with cap = 0 to size (returns n, 0 on error) */
SPT_API int spt_host_save_scene(const char *path, uint32_t n, const float *centers4, const float *radii, const uint8_t *materials,
                        const float *colors4, const float *fuzz, const float camera16[16]);
SPT_API uint32_t spt_host_load_scene(const char *path, uint32_t cap, float *centers4, float *radii, uint8_t *materials, float *colors4,
                             float *fuzz, float camera16[16]);
SPT_API int spt_host_save_pfm(const char *path, uint32_t width, uint32_t height, const float *accum4);
SPT_API int spt_host_save_hdr(const char *path, uint32_t width, uint32_t height, const float *accum4);
SPT_API int spt_host_save_accum(const char *path, uint32_t width, uint32_t height, uint32_t samples_done, uint32_t depth, uint64_t seed,
                        const float *accum4);
/* returns 1 and fills samples_done / depth / seed / accum4 (width*height*4) when the file matches the image size */
SPT_API int spt_host_load_accum(const char *path, uint32_t width, uint32_t height, uint32_t *samples_done, uint32_t *depth,
                        uint64_t *seed, float *accum4);

#ifdef __cplusplus
}
#endif
#endif /* SPT_HOST_H */
