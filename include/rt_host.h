/*
 * rt_host.h — the reference's host surface, kept source-compatible, on top of the
 * C ABI in rtb200.h.  A caller written against the reference headers
 * (gpu/headers/{parser,parse_obj,scene,printer,raytracer}.h) compiles against this
 * header unchanged: same function names, argument meaning and error behaviour.
 */
#ifndef RT_HOST_H
#define RT_HOST_H

#include <stdio.h>
#include <cstddef>
#include "rtb200.h"

/* gpu/headers/colors.h:14-19 — framebuffer element */
#ifndef COLORS_H
#define COLORS_H
struct color { uint8_t r, g, b, a; };
#endif

/* gpu/headers/scene.h:40-61 */
struct scene empty_scene();
struct object *add_object_to_scene(struct scene *scene, uint32_t nb_triangles);
vector3 *get_vertex(const struct triangles_layout triangles, uint32_t triangle_index);
vector3 *get_normal(const struct triangles_layout triangles, uint32_t triangle_index);
/* Frees everything empty_scene/add_object_to_scene/parser allocated (the reference
 * leaks; rt.cpp has no cleanup). */
void free_scene(struct scene *scene);

/* gpu/headers/parser.h:18, gpu/headers/parse_obj.h:8-9.
 * errx(1, ...) on an unopenable file or unknown token, like the reference. */
struct scene parser(const char *path);
void parse_object(FILE *file, struct scene *scene);
vector3 create_vec(FILE *file);
/* Non-exiting variant for tests/bindings: returns 0 or -1 and fills err[errlen]. */
extern "C" int rtb_parse_svati(const char *path, struct scene *out, char *err, size_t errlen);
extern "C" void rtb_free_scene(struct scene *scene);
/* Extension: room for `triangles` triangles in total now, so that the add_object_to_scene
 * calls that follow never move the vertex / normal arrays (loaders that know the size). */
extern "C" int rtb_scene_reserve(struct scene *scene, size_t triangles);
/* Extension: 1 = bulk arrays allocated by this thread from now on are plain memory instead of
 * page-locked (page-locking costs ~1 s per GB; a one-shot loader does not earn it back). */
extern "C" void rtb_scene_mem_pageable(int on);

/* gpu/headers/printer.h:11-12 (PPM P3 text) */
FILE *open_output(const char *output, int width, int height);
void print_color(struct color color, FILE *file);

/* gpu/rt.cpp:14-52 — 8-bit RGBA, non-interlaced PNG (own encoder on zlib; libpng is
 * not in this image).  Returns 0 on success. */
extern "C" int rtb_write_png(const unsigned char *buffer, int width, int height,
                             ptrdiff_t stride, const char *filename);
inline void write_png(const std::byte *buffer, int width, int height, int stride,
                      const char *filename) {
  rtb_write_png(reinterpret_cast<const unsigned char *>(buffer), width, height, stride,
                filename);
}

/* gpu/headers/raytracer.h:6 — prints the CUDA error and exit(1)s on failure, like
 * abortError (gpu/raytracer.cu:21-29). */
void render(const scene &scene, char *buffer, int aliasing, std::ptrdiff_t stride,
            int pre_h, int pre_w);

/* Synthetic benchmark scene of BASELINE.json configs[4], built through empty_scene /
 * add_object_to_scene / get_vertex / get_normal: a grid x grid tessellated height field
 * (2*grid^2 triangles, 16 objects, ambient + directional + point light).  grid = 2237
 * gives 10 008 338 triangles.  width/height: camera resolution (not multiplied by aliasing). */
extern "C" int rtb_synth_terrain(struct scene *out, int grid, uint64_t seed, int width, int height);
/* Scene -> .svati text that parser() reads back to the identical memory image. */
extern "C" int rtb_write_svati(const struct scene *scene, const char *path, int width, int height);

/* The body of main() in gpu/rt.cpp:54-97 as a callable (used by the `rt` binary). */
extern "C" int rtb_cli_main(int argc, char **argv);

#endif /* RT_HOST_H */
