// PPM (P3, text) writer behind the reference's printer API
// (gpu/headers/printer.h:11-12; behaviour of cpu/printer.c:3-18: header
// "P3\nW H\n255\n", then "r g b " per pixel, errx(1) when the file cannot be opened).
#include <cerrno>
#include <cstring>
#include <err.h>
#include "rt_host.h"

FILE *open_output(const char *output, int width, int height) {
  FILE *f = fopen(output, "w+");
  if (!f) errx(1, "%s\n", strerror(errno));
  fprintf(f, "P3\n%d %d\n255\n", width, height);
  return f;
}

void print_color(struct color c, FILE *file) {
  fprintf(file, "%d %d %d ", (int)c.r, (int)c.g, (int)c.b);
}
