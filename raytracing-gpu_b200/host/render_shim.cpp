// render() with the reference's C++ signature (gpu/headers/raytracer.h:6) on top of the
// C ABI.  Keeps the observable behaviour of gpu/raytracer.cu:177-253: the two progress
// lines on stdout (:221, :234) and "message + exit(1)" on any device failure (:21-29).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "rt_host.h"

void render(const scene &scene, char *buffer, int aliasing, std::ptrdiff_t stride, int pre_h, int pre_w) {
  const int bsize = 16;   // the reference's launch geometry, printed only
  const int wi = (int)std::ceil((float)scene.camera.width / bsize);
  const int he = (int)std::ceil((float)scene.camera.height / bsize);
  printf("lancement. %i %i %i %i.\n", wi, he, scene.camera.width, scene.camera.height);
  // RT_GPUS=n: the same frame on n GPUs of this process (argv stays that of the reference)
  const char *g = getenv("RT_GPUS");
  const int gpus = g ? atoi(g) : 1;
  const int rc = gpus > 1 ? rtb_render_multi(&scene, reinterpret_cast<unsigned char *>(buffer), aliasing, stride, pre_h,
                                             pre_w, gpus)
                          : rtb_render(&scene, reinterpret_cast<unsigned char *>(buffer), aliasing, stride, pre_h, pre_w);
  printf("done..\n");
  if (rc != RTB_OK) {
    fprintf(stderr, "[error] Computation Error (rtb200 %d): %s\n", rc, rtb_last_error());
    std::exit(1);
  }
}
