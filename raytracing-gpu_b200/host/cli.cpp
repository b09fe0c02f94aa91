// Body of the `rt` command (gpu/rt.cpp:54-97): rt file.svati output.png
// Same argv contract, banner, aliasing factor and exit codes.  Extras that do not change
// the contract: an output name ending in .ppm goes through the reference's printer API
// (P3 text) instead of PNG; RT_STATS=1 prints one JSON line with timings on stderr.
#include <err.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include "rt_host.h"

static bool ends_with(const char *s, const char *suffix) {
  const size_t n = strlen(s), m = strlen(suffix);
  return n >= m && strcmp(s + n - m, suffix) == 0;
}

extern "C" int rtb_cli_main(int argc, char **argv) {
  if (argc != 3) errx(1, "usage: %s file.svati output.png", argv[0]);
  std::cout << "Using structure of arrays (SOA) layout" << std::endl;
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  struct scene scene = parser(argv[1]);
  const auto t1 = clk::now();

  const int aliasing = 3;
  const int stride = scene.camera.width * (int)sizeof(struct color);
  auto buffer = std::make_unique<std::byte[]>((size_t)scene.camera.height * stride);
  const int pre_h = scene.camera.height, pre_w = scene.camera.width;
  scene.camera.width *= aliasing;
  scene.camera.height *= aliasing;
  render(scene, reinterpret_cast<char *>(buffer.get()), aliasing, stride, pre_h, pre_w);
  const auto t2 = clk::now();

  if (ends_with(argv[2], ".ppm")) {
    FILE *out = open_output(argv[2], pre_w, pre_h);
    const struct color *px = reinterpret_cast<const struct color *>(buffer.get());
    for (int y = 0; y < pre_h; ++y) {
      for (int x = 0; x < pre_w; ++x) print_color(px[(size_t)y * pre_w + x], out);
      fputc('\n', out);
    }
    fclose(out);
  } else {
    write_png(buffer.get(), pre_w, pre_h, stride, argv[2]);
  }
  const auto t3 = clk::now();
  const char *st = getenv("RT_STATS");
  if (st && *st == '1') {
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    fprintf(stderr, "{\"parse_ms\": %.3f, \"render_ms\": %.3f, \"encode_ms\": %.3f, \"triangles\": %zu, \"width\": %d, \"height\": %d}\n",
            ms(t0, t1), ms(t1, t2), ms(t2, t3), scene.triangle_count, pre_w, pre_h);
  }
  free_scene(&scene);
  return 0;
}
