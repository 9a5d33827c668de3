/*
 * host/ppm_out.c -- byte-exact P3 writer (format of C/ppm.c:168-184):
 *     "P3\n" "<w> <h> \n" "<maxColor>\n"   then one decimal per line, R G B per pixel.
 * (the space before the second newline is the reference's.)  The reference calls
 * fprintf once per pixel (~0.2 s per Mpixel); this one formats through a 256-entry
 * table into a 4 MiB buffer, and indexes with size_t so 32768^2 images work.
 */
#include <stdlib.h>
#include <string.h>
#include "ppm_out.h"

void ppm_WriteOutP3(PPMFormat inData, FILE *outFile) {
  char lut[256][8];   /* "255\n" + NUL needs 5 */
  uint8_t lut_len[256];
  for (int v = 0; v < 256; v++) lut_len[v] = (uint8_t)sprintf(lut[v], "%d\n", v);

  fprintf(outFile, "P3\n%d %d \n%u\n", inData.width, inData.height, inData.maxColor);

  const size_t total = (size_t)inData.width * (size_t)inData.height * 3u;
  const size_t cap = (size_t)4 << 20;
  char *buf = (char *)malloc(cap + 8);
  if (!buf) { fprintf(stderr, "Error: out of memory in ppm_WriteOutP3\n"); exit(1); }
  size_t fill = 0;
  const uint8_t *src = inData.pixmap;
  for (size_t i = 0; i < total; i++) {
    const uint8_t v = src[i];
    memcpy(buf + fill, lut[v], 4);
    fill += lut_len[v];
    if (fill >= cap) { fwrite(buf, 1, fill, outFile); fill = 0; }
  }
  if (fill) fwrite(buf, 1, fill, outFile);
  free(buf);
}

void ppm_WriteOutP6(PPMFormat inData, FILE *outFile) {
  fprintf(outFile, "P6\n%d %d\n%u\n", inData.width, inData.height, inData.maxColor);
  fwrite(inData.pixmap, 1, (size_t)inData.width * (size_t)inData.height * 3u, outFile);
}

float ppm_clamp(float value, float lower_bound, float upper_bound) {
  if (value > upper_bound) value = upper_bound;
  if (value < lower_bound) value = lower_bound;
  return value;
}
